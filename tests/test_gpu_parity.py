"""GPU parity tests: every CUDA kernel, called through the C ABI (b200seg.ops -> libtss_b200.so), against
the CPU oracle on the same seeded inputs and against the reference's golden fixtures.

Tolerances (BASELINE.json north_star): labels / confusion matrices bit-exact except at pixels whose top-2
margin or threshold gap is < 1e-6; prototypes, distances, losses within 1e-5 relative (fp32 accumulation).
For vectors "relative" is norm-wise: |a-b| <= 1e-5 * max|b| (element-wise relative error is meaningless at
zero crossings of a mean of N(0,1) features).
"""
import numpy as np
import pytest
import torch

import inputs as gi
from oracle import label_path as O

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def weight_rtol(d_ref, tau, rtol_d=RTOL):
    """Propagated bound for w = softmax(-tau (d - min d)): dw_c / w_c = -tau (dd_c - sum_j w_j dd_j), hence
    |dw / w| <= 2 tau max|dd| <= 2 tau rtol_d max(d) when the distances are within rtol_d relative.
    north_star's 1e-5 is a bound on the distances; the weights inherit it through this factor (25 at K = 256)."""
    return 2.0 * tau * rtol_d * float(np.max(d_ref))


def rectify_mismatches_allowed(rect, bad, train_thred, eta):
    """argmax / ignore decisions of the rectify block may differ only where the reference's own normalised
    rectified probabilities have a top-2 margin or a threshold gap inside the propagated weight error eta + 1e-6"""
    srt = np.sort(rect, axis=1)
    top, second = srt[:, -1:], srt[:, -2:-1]
    ok = ((top - second) < 1e-6 + 2 * eta * top) | (np.abs(top - train_thred) < 1e-6 + eta * top)
    return bool(ok[bad].all())


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from b200seg import ops as _ops
    _ops.lib()
    return _ops


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def G(a):
    return T(a).cuda()


def close_vec(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max()
    assert err <= rtol * scale, "max err %.3e vs scale %.3e (rel %.3e)" % (err, scale, err / scale)


def top2_margin(p, axis=1):
    s = np.sort(p, axis=axis)
    return np.take(s, -1, axis=axis) - np.take(s, -2, axis=axis)


def fp16_ulp_diff(a, b):
    ia = a.view(np.int16).astype(np.int32)
    ib = b.view(np.int16).astype(np.int32)
    return np.abs(ia - ib)


# ------------------------------------------------------------------------------------------- A14
@pytest.mark.parametrize("c", [13, 19])
@pytest.mark.parametrize("dtype", [torch.uint8, torch.int32, torch.int64])
def test_confusion_golden(ops, golden, c, dtype):
    pred, gt = gi.hist_inputs(seed=601 + c, c=c)
    if dtype == torch.uint8:
        gt = np.where(gt < 0, 255, gt)  # uint8 cannot hold -1: both are "ignored"
    h, bad = ops.confusion_matrix(G(pred).to(dtype), G(gt).to(dtype), c)
    assert int(bad.item()) == 0
    assert np.array_equal(h.cpu().numpy(), golden["hist_c%d" % c])


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("n", [0, 1, 15, 16 * 32 * 15 + 5, 3_000_001])
def test_confusion_sizes_and_variants(ops, n, variant):
    r = gi.rs(700 + n % 97)
    c = 19
    gt = r.randint(0, 21, size=n).astype(np.uint8)
    gt[r.uniform(size=n) < 0.01] = 255
    pred = np.where(r.uniform(size=n) < 0.7, np.minimum(gt, c - 1), r.randint(0, c, size=n)).astype(np.uint8)
    h, bad = ops.confusion_matrix(G(pred), G(gt), c, variant=ops.HIST_PACKED if variant else ops.HIST_ATOMIC)
    ref = O.fast_hist(pred.astype(np.int64), np.where(gt == 255, -1, gt).astype(np.int64), c) if n else np.zeros((c, c), np.int64)
    assert np.array_equal(h.cpu().numpy(), ref)
    assert int(bad.item()) == 0
    # accumulation (+=) and a checksum of checksums: total == number of valid pixels
    h2, _ = ops.confusion_matrix(G(pred), G(gt), c, hist=h.clone())
    assert np.array_equal(h2.cpu().numpy(), 2 * ref)
    assert int(h.sum().item()) == int((gt < c).sum())


@pytest.mark.parametrize("c", [2, 13, 15])
@pytest.mark.parametrize("n", [512 * 148 * 8, 5_000_003])
def test_confusion_large_runs_and_bad_values(ops, c, n):
    """large n: out-of-range values in both arrays, a tail that is not a multiple of 512, long runs of one key
    (far more than 255 increments of one packed 8-bit counter between flushes); both kernels agree"""
    r = gi.rs(800 + c)
    gt = r.randint(0, c + 3, size=n).astype(np.uint8)
    gt[r.uniform(size=n) < 0.02] = 255
    gt[1000:300000] = min(1, c - 1)                       # one key for ~300k consecutive pixels
    pred = np.where(r.uniform(size=n) < 0.7, np.minimum(gt, c - 1), r.randint(0, c, size=n)).astype(np.uint8)
    pred[r.uniform(size=n) < 0.001] = 250                  # out-of-range predictions -> bad counter
    h, bad = ops.confusion_matrix(G(pred), G(gt), c)
    ok = gt < c
    okp = ok & (pred < c)
    ref = np.bincount(gt[okp].astype(np.int64) * c + pred[okp], minlength=c * c).reshape(c, c)
    assert np.array_equal(h.cpu().numpy(), ref)
    assert int(bad.item()) == int((ok & (pred >= c)).sum())
    h1, bad1 = ops.confusion_matrix(G(pred), G(gt), c, variant=ops.HIST_ATOMIC)
    assert torch.equal(h, h1) and torch.equal(bad, bad1)


def test_confusion_bad_pred_reported(ops):
    gt = np.array([0, 1, 2, 3] * 5000, np.uint8)
    pred = gt.copy()
    pred[7] = 200
    h, bad = ops.confusion_matrix(G(pred), G(gt), 13)
    assert int(bad.item()) == 1 and int(h.sum().item()) == gt.size - 1


def test_confusion_rejects_cpu_tensors(ops):
    with pytest.raises(ops.TssError):
        ops.confusion_matrix(torch.zeros(4, dtype=torch.uint8), torch.zeros(4, dtype=torch.uint8).cuda(), 13)


# ------------------------------------------------------------------------------------------- A1-A4
def _check_labels(lab, conf32, ref_lab, ref_prob, what):
    """labels exact except where the reference's own top-2 margin is < 1e-6"""
    lab, ref_lab = np.asarray(lab, np.int64), np.asarray(ref_lab, np.int64)
    bad = lab != ref_lab
    if bad.any():
        margin = top2_margin(ref_prob, 1)
        assert (margin[bad] < 1e-6).all(), "%s: %d label mismatches outside the 1e-6 margin" % (what, int(bad.sum()))


@pytest.mark.parametrize("name", list(gi.PL_CASES))
def test_pseudo_label_golden_branches(ops, golden, name):
    n, c, h, w, H, W, _ = gi.PL_CASES[name]
    logits, flip_logits = gi.pl_logits(name)
    # soft branch
    prob = ops.softmax_nchw(G(logits)).cpu().numpy()
    np.testing.assert_allclose(prob, golden[name + "_soft_prob"], rtol=2e-6, atol=1e-9)
    # hard branch, no flip: softmax + max at network resolution
    out = ops.pseudo_label(G(logits), None, ops.PL_UP_SOFTMAX, want_conf32=True)
    ref_prob = golden[name + "_soft_prob"]
    _check_labels(out.label.cpu().numpy(), None, golden[name + "_hard_label"], ref_prob, "hard")
    # the fp32 confidence is within ~3e-7 relative of the reference's (ex2.approx + refined rcp), so its fp16
    # storage cast lands on the neighbouring fp16 value for a fraction of about 2 * 3e-7 / 2^-11 = 0.12 % of pixels
    assert fp16_ulp_diff(out.conf16.cpu().numpy(), golden[name + "_hard_conf"]).max() <= 1
    assert (fp16_ulp_diff(out.conf16.cpu().numpy(), golden[name + "_hard_conf"]) > 0).mean() < 0.005
    # flip branch exactly as committed: class-axis "fliplr", second softmax
    pa, pb = ops.softmax_nchw(G(logits)), ops.softmax_nchw(G(flip_logits))
    out = ops.pseudo_label(pa, (H, W), ops.PL_PROB_UP_SOFTMAX, src_flip=pb, flip=ops.FLIP_CLASS, want_conf32=True)
    a = O.upsample_bilinear_ac(O.pseudo_label_soft(T(logits)), (H, W))
    b = O.upsample_bilinear_ac(O.pseudo_label_soft(T(flip_logits)), (H, W))
    avg = ((a + torch.flip(b, dims=[1])) / 2).numpy()
    ref2 = torch.softmax(T(avg), 1).numpy()
    _check_labels(out.label.cpu().numpy(), None, golden[name + "_flip_label"], ref2, "flip")
    assert fp16_ulp_diff(out.conf16.cpu().numpy(), golden[name + "_flip_conf"]).max() <= 1
    # evaluation branch: upsample + argmax
    out = ops.pseudo_label(G(logits), (H, W), ops.PL_UP_ARGMAX, want_conf16=False)
    up = O.upsample_bilinear_ac(T(logits), (H, W)).numpy()
    _check_labels(out.label.cpu().numpy(), None, golden[name + "_eval_pred"], up, "eval")
    # plain upsample kernel against the golden planes
    mine = ops.upsample_bilinear(G(logits), (H, W)).cpu().numpy()
    assert np.abs(mine[:, :3] - golden[name + "_up_logits_c0to2"]).max() <= 2e-6


@pytest.mark.parametrize("shape", [(2, 13, 41, 81, 320, 640), (1, 19, 17, 33, 128, 256), (1, 5, 9, 9, 9, 9),
                                   (1, 13, 11, 23, 37, 91), (1, 32, 8, 8, 64, 64)])
def test_pseudo_label_upsample_softmax_vs_oracle(ops, shape):
    n, c, h, w, H, W = shape
    logits = gi.randn(900 + c + H, n, c, h, w, scale=3.0)
    thr = np.linspace(0.2, 0.5, c).astype(np.float32)
    keep = list(range(0, c, 2))
    out = ops.pseudo_label(G(logits), (H, W), ops.PL_UP_SOFTMAX, thr=thr, keep=keep, ignore_index=255, want_conf32=True)
    up = O.upsample_bilinear_ac(T(logits), (H, W))
    prob = torch.softmax(up, 1).numpy()
    conf_ref, lab_ref = O.pseudo_label_upsampled(T(logits), (H, W))
    lab = out.label.cpu().numpy()
    _check_labels(lab, None, lab_ref.numpy(), prob, "up_softmax")
    same = lab == lab_ref.numpy()
    c32 = out.conf32.cpu().numpy()
    assert np.abs(c32 - conf_ref.numpy())[same].max() <= 2e-6
    c16 = out.conf16.cpu().numpy()
    assert np.array_equal(c16, c32.astype(np.float16))  # storage cast is round-to-nearest-even
    # threshold block applied to OUR (label, fp16 conf) must equal the oracle's threshold on the same pair
    thr_ref = O.threshold_hard_label(lab.reshape(-1), c16.reshape(-1), thr, 255, keep=keep).reshape(lab.shape)
    assert np.array_equal(out.label_thr.cpu().numpy(), thr_ref)
    # and the standalone threshold kernel agrees
    again = ops.threshold_labels(out.label, out.conf16, c, thr=thr, keep=keep, ignore_index=255)
    assert np.array_equal(again.cpu().numpy(), thr_ref)


def test_pseudo_label_direct_equals_tiled(ops):
    """full-resolution input: the 128-bit direct kernel and the tiled kernel give the same bits"""
    logits = gi.randn(950, 2, 13, 64, 128, scale=3.0)
    a = ops.pseudo_label(G(logits), None, ops.PL_UP_SOFTMAX, want_conf32=True)
    b = ops.pseudo_label(G(logits), None, ops.PL_UP_SOFTMAX, want_conf32=True, flags=ops.pl_flags(force_tile=True))
    assert torch.equal(a.label, b.label) and torch.equal(a.conf32, b.conf32)


def test_pseudo_label_width_flip(ops):
    logits, flip_logits = gi.pl_logits("pl_small")
    H, W = 64, 128
    pa, pb = ops.softmax_nchw(G(logits)), ops.softmax_nchw(G(flip_logits))
    out = ops.pseudo_label(pa, (H, W), ops.PL_PROB_UP, src_flip=pb, flip=ops.FLIP_WIDTH, want_conf32=True)
    a = O.upsample_bilinear_ac(O.pseudo_label_soft(T(logits)), (H, W))
    b = O.upsample_bilinear_ac(O.pseudo_label_soft(T(flip_logits)), (H, W))
    avg = ((a + torch.flip(b, dims=[3])) / 2)
    conf, lab = avg.max(1)
    _check_labels(out.label.cpu().numpy(), None, lab.numpy(), avg.numpy(), "wflip")
    same = out.label.cpu().numpy() == lab.numpy()
    assert np.abs(out.conf32.cpu().numpy() - conf.numpy())[same].max() <= 2e-6


def test_threshold_block_golden(ops, golden):
    label, conf = gi.threshold_inputs()
    kept = ops.threshold_labels(G(label), G(conf.astype(np.float16)), 13, thr=0.3, keep=(5, 11), ignore_index=12)
    out = ops.resize_nearest_u8(kept, (320, 700), rule="pil")
    assert np.array_equal(out.cpu().numpy(), golden["thr_label_hard_700x320"])


def test_nearest_torch_rule(ops):
    lab = gi.rs(77).randint(0, 13, size=(2, 1, 64, 128)).astype(np.uint8)
    ref = O.resize_nearest_torch(T(lab.astype(np.int64)), (16, 32)).numpy()
    out = ops.resize_nearest_u8(G(lab), (16, 32), rule="torch")
    assert np.array_equal(out.cpu().numpy().astype(np.int64), ref)


# ------------------------------------------------------------------------------------------- A5/A6
def _means_to_lists(m):
    valid = m.valid.cpu().numpy().astype(bool)
    vec = m.vectors.cpu().numpy()
    ids, out = [], []
    for n in range(valid.shape[0]):
        for t in range(valid.shape[1]):
            if valid[n, t]:
                ids.append(t)
                out.append(vec[n, t])
    return np.array(ids, np.int64), (np.stack(out) if out else np.zeros((0, vec.shape[2]), np.float32))


def test_class_means_golden(ops, golden):
    feat, logits, labels_val = gi.proto_inputs()
    m = ops.class_means(G(feat), G(logits), 13)
    ids, vec = _means_to_lists(m)
    assert np.array_equal(ids, golden["cf_ids"])
    for a, b in zip(vec, golden["cf_vectors"]):
        close_vec(a, b)
    m = ops.class_means(G(feat), G(logits), 13, labels_val=G(labels_val))
    ids, vec = _means_to_lists(m)
    assert np.array_equal(ids, golden["cf_val_ids"])
    for a, b in zip(vec, golden["cf_val_vectors"]):
        close_vec(a, b)
    for tag, th in (("none", None), ("t05", 0.5)):
        m = ops.class_means(G(feat), G(logits), 13, overflow_channel=12, conf_thresh=th, use_conf_gate=True)
        ids, vec = _means_to_lists(m)
        assert np.array_equal(ids, golden["st_mv_%s_ids" % tag])
        for a, b in zip(vec, golden["st_mv_%s_vectors" % tag]):
            close_vec(a, b)


def test_prototype_update_golden(ops, golden):
    feat, logits, _ = gi.proto_inputs()
    protos = torch.zeros(13, 256, device="cuda")
    num = torch.zeros(13, device="cuda")
    ops.prototype_update(protos, num, ops.class_means(G(feat), G(logits), 13), ops.UPDATE_MEAN, True, 0.9999)
    feat2 = gi.randn(302, *feat.shape)
    ops.prototype_update(protos, num, ops.class_means(G(feat2), G(logits), 13), ops.UPDATE_MEAN, True, 0.9999)
    assert np.array_equal(num.cpu().numpy(), golden["cf_objective_num"])
    for a, b in zip(protos.cpu().numpy(), golden["cf_objective_vectors"]):
        close_vec(a, b)
    # moving average from loaded prototypes (self_training.py:125-126, start_mean=False)
    m = ops.class_means(G(feat), G(logits), 13, overflow_channel=12, use_conf_gate=True)
    protos = G(gi.selftrain_protos()).clone()
    num = torch.zeros(13, device="cuda")
    ops.prototype_update(protos, num, m, ops.UPDATE_MOVING_AVERAGE, False, 1e-4)
    assert np.array_equal(num.cpu().numpy(), golden["st_ma_num"])
    np.testing.assert_allclose(protos.cpu().numpy(), golden["st_ma_vectors"], rtol=1e-6, atol=1e-7)
    # default start_mean=True: num < 100 forces 'mean'
    protos = torch.zeros(13, 256, device="cuda")
    num = torch.zeros(13, device="cuda")
    ops.prototype_update(protos, num, m, ops.UPDATE_MOVING_AVERAGE, True, 1e-4)
    assert np.array_equal(num.cpu().numpy(), golden["st_sm_num"])
    for a, b in zip(protos.cpu().numpy(), golden["st_sm_vectors"]):
        close_vec(a, b)


@pytest.mark.parametrize("shape", [(2, 256, 13, 41, 81), (1, 2048, 19, 33, 65), (3, 64, 7, 5, 7), (1, 32, 13, 1, 700)])
def test_class_means_vs_oracle_shapes(ops, shape):
    n, k, c, h, w = shape
    feat = gi.randn(1000 + k + h, n, k, h, w)
    logits = gi.randn(1001 + k + h, n, c, h, w, scale=3.0)
    m = ops.class_means(G(feat), G(logits), c)
    ids, vec = _means_to_lists(m)
    vr, ir = O.class_mean_vectors(T(feat), T(logits), c)
    assert np.array_equal(ids, np.array(ir, np.int64))
    for a, b in zip(vec, vr):
        close_vec(a, b.reshape(-1).numpy())
    # segmented-sum linearity: the per-class sums add up to the plain sum over pixels of every channel
    tot = m.sums.sum(1).cpu().numpy()
    close_vec(tot, feat.reshape(n, k, -1).astype(np.float64).sum(-1), rtol=2e-5)
    assert int(m.count_all.sum().item()) == n * h * w


def test_prototype_update_partial_matches_sequential(ops):
    """closed-form sharded update (2 'ranks' emulated on one GPU) == sequential kernel == oracle"""
    n, k, c, h, w = 6, 64, 13, 9, 17
    feat = gi.randn(1100, n, k, h, w)
    logits = gi.randn(1101, n, c, h, w, scale=3.0)
    for mode, start_mean, mom in ((ops.UPDATE_MEAN, True, 0.9999), (ops.UPDATE_MOVING_AVERAGE, False, 1e-2)):
        protos0 = G(gi.randn(1102, c, k))
        num0 = torch.full((c,), 2.0, device="cuda")
        full = ops.class_means(G(feat), G(logits), c)
        seq_p, seq_n = protos0.clone(), num0.clone()
        ops.prototype_update(seq_p, seq_n, full, mode, start_mean, mom)
        halves = [ops.class_means(G(feat[:3]), G(logits[:3]), c), ops.class_means(G(feat[3:]), G(logits[3:]), c)]
        counts = [ops.prototype_update_partial(hm, num0, mode, start_mean, mom, None, None) for hm in halves]
        total = (counts[0] + counts[1]).to(torch.int32)
        firsts = [torch.zeros_like(counts[0]), counts[0].clone()]
        acc = torch.zeros_like(protos0)
        for hm, fi in zip(halves, firsts):
            part, scale0, num_after = ops.prototype_update_partial(hm, num0, mode, start_mean, mom, fi, total)
            acc += part
        new_p = scale0[:, None] * protos0 + acc
        assert np.array_equal(num_after.cpu().numpy(), seq_n.cpu().numpy())
        for a, b in zip(new_p.cpu().numpy(), seq_p.cpu().numpy()):
            close_vec(a, b)


def test_finalize_scatter_writes_every_destination(ops):
    """tss_proto_finalize_scatter: the mean vectors land in vectors_out and in every extra destination (the peers'
    gather slots at N > 1; plain device buffers here), identical to tss_proto_finalize"""
    n, k, c, h, w = 3, 64, 13, 19, 23
    feat, logits = G(gi.randn(1500, n, k, h, w)), G(gi.randn(1501, n, c, h, w, scale=3.0))
    ref = ops.class_means(feat, logits, c)
    out = torch.full((n, c, k), -7.0, device="cuda")
    extra = [torch.full((n, c, k), -7.0, device="cuda") for _ in range(3)]
    out_s = torch.full((n, c), -7.0, device="cuda")
    extra_s = [torch.full((n, c), -7.0, device="cuda") for _ in range(3)]
    got = ops.class_means(feat, logits, c, scatter=(out, out_s, [e.data_ptr() for e in extra], [e.data_ptr() for e in extra_s]))
    assert got.vectors.data_ptr() == out.data_ptr()
    for t in extra:                       # every destination of ONE launch holds the same bits
        assert torch.equal(t, out)
    for t in extra_s:
        assert torch.equal(t, out_s)
    # two runs of the accumulate kernel differ in the last bits (fp32 RED.ADD order is not fixed): 1e-6, not equality
    close_vec(out.cpu().numpy(), ref.vectors.cpu().numpy(), rtol=2e-6)
    close_vec(out_s.cpu().numpy(), ref.vector_sum.cpu().numpy(), rtol=2e-5)   # a sum over K of values of both signs
    assert torch.equal(got.valid, ref.valid) and torch.equal(got.count_gated, ref.count_gated)


# ------------------------------------------------------------------------------------------- A7/A8/A10
def test_distance_weight_golden(ops, golden):
    protos = G(gi.selftrain_protos())
    feat = G(gi.selftrain_feat())
    d, _, _ = ops.prototype_distance(feat, protos)
    np.testing.assert_allclose(d.cpu().numpy(), golden["st_distance"], rtol=RTOL)
    f2w = O.full2weak(T(gi.full2weak_feat()), gi.WEAK_PARAMS)
    for form in (0, ops.dist_form(algo=ops.DIST_DIRECT), ops.dist_form(algo=ops.DIST_EXPANDED)):
        d2, wgt, _ = ops.prototype_distance(f2w.cuda(), protos, temperature=0.5, want_dist=True, want_weight=True, form=form)
        d_ref = O.prototype_distance(f2w, protos.cpu()).numpy()
        np.testing.assert_allclose(d2.cpu().numpy(), d_ref, rtol=RTOL)
        np.testing.assert_allclose(wgt.cpu().numpy(), golden["st_weight"], rtol=weight_rtol(d_ref, 0.5), atol=1e-30)
    # the same weights straight from the quarter-resolution map through the fused sampling kernel (a 1:1 first stage)
    _, wgt, _ = ops.prototype_distance_sampled(G(gi.full2weak_feat()), (20, 40), gi.WEAK_PARAMS, protos, temperature=0.5,
                                               want_weight=True)
    np.testing.assert_allclose(wgt.cpu().numpy(), golden["st_weight"], rtol=weight_rtol(d_ref, 0.5), atol=1e-30)


@pytest.mark.parametrize("shape", [(1, 2048, 19, 17, 33), (2, 256, 13, 41, 81), (1, 32, 4, 3, 5), (1, 96, 20, 8, 8)])
def test_distance_vs_oracle_shapes(ops, shape):
    n, k, c, h, w = shape
    feat = gi.randn(1300 + k, n, k, h, w)
    protos = gi.randn(1301 + k, c, k)
    soft = gi.softmax_np(gi.randn(1302 + k, n, c, h, w, scale=3.0), 1)
    hard = gi.rs(1303).choice([0, 1, 255, 255], size=(n, 1, h, w)).astype(np.int64)
    d, wgt, lab = ops.prototype_distance(G(feat), G(protos), temperature=0.5, want_dist=True, want_weight=True,
                                         soft=G(soft), hard=G(hard), train_thred=0.3, ignore_index=255)
    d_ref = O.prototype_distance(T(feat), T(protos))
    np.testing.assert_allclose(d.cpu().numpy(), d_ref.numpy(), rtol=RTOL)
    rel_d = float(np.max(np.abs(d.cpu().numpy() - d_ref.numpy()) / d_ref.numpy()))
    w_ref = O.prototype_weight(T(feat), T(protos), 0.5)
    np.testing.assert_allclose(wgt.cpu().numpy(), w_ref.numpy(), rtol=weight_rtol(d_ref.numpy(), 0.5), atol=1e-30)
    ref_lab, rect = O.rectify_labels(w_ref, T(soft), T(hard), 0.3, 255)
    bad = (lab.cpu() != ref_lab).numpy()
    if bad.any():   # decisions differ only inside the weight error propagated from the MEASURED distance error
        eta = 2 * 0.5 * max(rel_d, 1e-7) * float(d_ref.max())
        assert rectify_mismatches_allowed(rect.numpy(), bad, 0.3, eta)
        assert bad.mean() < 1e-3


def test_distance_expanded_form_near_prototypes(ops):
    """The expanded form |f|^2 + |p|^2 - 2 f.p cancels when a feature sits on a prototype: features
    = prototype + 1e-3 noise on a third of the pixels, = 0.9 prototype on a few, random elsewhere.  Every (pixel,
    class) inside the cancellation region must come out of the exact recomputation: 1e-5 relative everywhere."""
    n, k, c, h, w = 2, 2048, 19, 17, 33
    r = gi.rs(1700)
    protos = gi.randn(1701, c, k)
    feat = gi.randn(1702, n, k, h, w)
    near = r.uniform(size=(n, h, w)) < 0.33
    cls = r.randint(0, c, size=(n, h, w))
    noise = (1e-3 * r.standard_normal((n, k, h, w))).astype(np.float32)
    on_proto = np.transpose(protos[cls], (0, 3, 1, 2))             # [n, k, h, w]
    feat = np.where(near[:, None], on_proto + noise, feat).astype(np.float32)
    scaled = r.uniform(size=(n, h, w)) < 0.05
    feat = np.where(scaled[:, None] & ~near[:, None], 0.9 * on_proto, feat).astype(np.float32)
    feat[0, :, 0, 0] = protos[3]                                    # exactly on a prototype: distance 0
    d_ref = O.prototype_distance(T(feat), T(protos)).numpy()
    assert d_ref[0, 3, 0, 0] == 0.0 and d_ref.min() == 0.0 and (d_ref < 0.1).mean() > 0.01
    for form in (ops.dist_form(8), ops.dist_form(algo=ops.DIST_EXPANDED), ops.dist_form(2, 1, 1, ops.DIST_EXPANDED),
                 ops.dist_form(4, 0, 2, ops.DIST_EXPANDED), ops.dist_form(2, 2, 4, ops.DIST_EXPANDED), 0):
        d, wgt, _ = ops.prototype_distance(G(feat), G(protos), temperature=0.5, want_dist=True, want_weight=True, form=form)
        np.testing.assert_allclose(d.cpu().numpy(), d_ref, rtol=RTOL, atol=1e-6, err_msg="form 0x%x" % form)
        w_ref = O.prototype_weight(T(feat), T(protos), 0.5).numpy()
        np.testing.assert_allclose(wgt.cpu().numpy(), w_ref, rtol=weight_rtol(d_ref, 0.5), atol=1e-30)


@pytest.mark.parametrize("variant", [0, 1, 2, 4, 8])
@pytest.mark.parametrize("split", [0, 1, 2, 4])
def test_distance_kernel_forms_agree(ops, variant, split):
    """every kernel form / K split / algorithm behind tss_proto_distance gives the oracle's distances (ragged plane:
    23*37 px); the form is a per-call argument (no process-wide tuning state)"""
    n, k, c, h, w = 2, 128, 19, 23, 37
    feat, protos = gi.randn(1400, n, k, h, w), gi.randn(1401, c, k)
    d_ref = O.prototype_distance(T(feat), T(protos)).numpy()
    for prefetch in (0, 1, 2):
        for algo in (ops.DIST_DIRECT, ops.DIST_EXPANDED):
            form = ops.dist_form(variant, prefetch, split, algo)
            d, _, _ = ops.prototype_distance(G(feat), G(protos), temperature=0.5, want_dist=True, form=form)
            np.testing.assert_allclose(d.cpu().numpy(), d_ref, rtol=RTOL, err_msg="form 0x%x" % form)


# ------------------------------------------------------------------------------------------- A11-A13
def test_ce_rce_golden(ops, golden):
    logits, label = gi.ce_inputs()
    lg = G(logits).requires_grad_(True)
    total, ce, rce, bad, _ = ops.ce_rce(lg, G(label), 12, 0.1, 1.0)
    total.backward()
    assert int(bad.item()) == 0
    np.testing.assert_allclose(ce.item(), golden["st_ce"], rtol=RTOL)
    np.testing.assert_allclose(rce.item(), golden["st_rce"], rtol=RTOL)
    close_vec(lg.grad.cpu().numpy(), golden["st_loss_grad"])
    # the whole-step target loss, from the reference's own recorded logits and rectified labels
    for tag in ("t0", "t06"):
        lt = ops.ce_rce(G(golden["step_%s_target_logits" % tag]), G(golden["step_%s_rectified_label" % tag]), 12, 0.1, 1.0)[0]
        np.testing.assert_allclose(lt.item(), golden["step_%s_loss_target" % tag], rtol=RTOL)


def test_semantic_consistency_and_focal_golden(ops, golden):
    logits, label = gi.ce_inputs(seed=502, c=13)
    tgt = G(label).clone()
    for cidx in (3, 5):
        tgt[tgt == cidx] = 255
    assert np.array_equal(tgt.cpu().numpy(), golden["sc_target_after"])
    lg = G(logits).requires_grad_(True)
    loss = ops.cross_entropy(lg, tgt, 255)
    loss.backward()
    np.testing.assert_allclose(loss.item(), golden["sc_loss"], rtol=RTOL)
    close_vec(lg.grad.cpu().numpy(), golden["sc_grad"])
    for tag, kw in (("g2", dict(gamma=2.0)), ("g2a", dict(gamma=2.0, alpha=[0.5 + 0.05 * i for i in range(13)])),
                    ("g0sum", dict(gamma=0.0, size_average=False))):
        lg = G(logits).requires_grad_(True)
        alpha = kw.pop("alpha", None)
        fl = ops.focal_loss(lg, G(label), alpha=None if alpha is None else torch.tensor(alpha).cuda(), **kw)
        fl.backward()
        np.testing.assert_allclose(fl.item(), golden["focal_%s_loss" % tag], rtol=RTOL)
        close_vec(lg.grad.cpu().numpy(), golden["focal_%s_grad" % tag])


def test_ce_all_ignored_and_bad_labels(ops):
    logits = G(gi.randn(1400, 1, 13, 8, 8))
    tgt = torch.full((1, 8, 8), 12, dtype=torch.int64, device="cuda")
    total, ce, rce, bad, _ = ops.ce_rce(logits, tgt, 12, 0.1, 1.0)
    assert torch.isnan(ce).item() and rce.item() == 0.0  # torch: mean over zero pixels is nan; rce has the +1e-6
    tgt[0, 0, 0] = 40
    assert int(ops.ce_rce(logits, tgt, 12, 0.1, 1.0)[3].item()) == 1
