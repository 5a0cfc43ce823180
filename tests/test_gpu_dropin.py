"""GPU parity of the drop-in mirror (b200seg.dropin) and of the later kernels (full2weak, one-hot, upsample
adjoint, single-vector update, sharded prototype form) against the CPU oracle, on seeded inputs."""
import types

import numpy as np
import pytest
import torch

from oracle import label_path as O

pytestmark = pytest.mark.gpu
RTOL = 1e-5  # north_star: prototypes, distances and loss within 1e-5 relative (fp32 accumulation)


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import dropin, ops, sharding
    return types.SimpleNamespace(dropin=dropin, ops=ops, sharding=sharding)


def rs(seed):
    return np.random.RandomState(seed)


def G(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def close(a, b, rtol=RTOL, atol=1e-6):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    assert np.abs(a - b).max() <= rtol * scale + atol, (np.abs(a - b).max(), scale)


def weight_rtol(d_ref, tau=0.5, rtol_d=RTOL):
    """|dw / w| <= 2 tau max|dd| for w = softmax(-tau (d - min d)); distances within rtol_d relative"""
    return 2.0 * tau * rtol_d * float(np.max(d_ref))


def _args(**kw):
    base = dict(num_classes=13, ignore_index=12, proto_temperature=0.5, proto_momentum=1e-4, train_thred=0.0,
                rce=True, rce_alpha=0.1, rce_beta=1.0, proto_rectify=True, ema=True, moving_prototype=True)
    base.update(kw)
    return types.SimpleNamespace(**base)


WEAK = {"RandomSized": [torch.tensor([563, 704]), torch.tensor([1126, 1408])],
        "RandomCrop": [torch.tensor([13, 101]), torch.tensor([269, 357]), torch.tensor([57, 300]), torch.tensor([569, 812])],
        "RandomHorizontallyFlip": [torch.tensor(True), torch.tensor(False)]}


# ------------------------------------------------------------------------------------------------
def test_class_features_lists_and_single_updates(mods):
    r = rs(11)
    feat = r.standard_normal((3, 256, 9, 17)).astype(np.float32)
    logits = (3 * r.standard_normal((3, 13, 9, 17))).astype(np.float32)
    logits[0, 2] = -50.0
    logits[0, 2, 0, :4] = 50.0                                   # 4 px only: the "< 10" skip
    cf = mods.dropin.Class_Features("cuda", 13)
    vecs, ids = cf.calculate_mean_vector(G(feat), G(logits))
    rvecs, rids = O.class_mean_vectors(torch.from_numpy(feat), torch.from_numpy(logits), 13)
    assert ids == rids
    bank = O.PrototypeBank(13, 256)
    for t, v, rv in zip(ids, vecs, rvecs):
        assert tuple(v.shape) == (256, 1, 1)
        close(v.cpu().numpy(), rv.numpy())
        cf.update_objective_SingleVector(t, v.cpu(), "mean")     # the reference passes a CPU vector (:75)
        bank.update(t, rv, "mean")
    close(cf.objective_vectors.cpu().numpy(), bank.vectors.numpy())
    assert torch.equal(cf.objective_vectors_num.cpu(), bank.num)
    # an all-zero vector is skipped, moving average after num >= 100 follows the reference formula
    cf.update_objective_SingleVector(0, torch.zeros(256, 1, 1), "mean")
    assert torch.equal(cf.objective_vectors_num.cpu(), bank.num)
    cf.objective_vectors_num[1] = 150.0
    bank.num[1] = 150.0
    v = torch.from_numpy(r.standard_normal((256, 1, 1)).astype(np.float32))
    cf.update_objective_SingleVector(1, v, "moving_average")
    bank.update(1, v, "moving_average")
    close(cf.objective_vectors.cpu().numpy(), bank.vectors.numpy())
    with pytest.raises(NotImplementedError):
        cf.update_objective_SingleVector(1, v, "median")
    # batched fast path == list path
    cf2 = mods.dropin.Class_Features("cuda", 13)
    cf2.update_from_batch(G(feat), G(logits))
    bank2 = O.PrototypeBank(13, 256)
    for t, rv in zip(rids, rvecs):
        bank2.update(t, rv, "mean")
    close(cf2.objective_vectors.cpu().numpy(), bank2.vectors.numpy())


def test_process_label_one_hot(mods):
    r = rs(12)
    lab = r.randint(0, 15, size=(2, 1, 7, 9)).astype(np.float32)   # 13, 14 overflow
    cf = mods.dropin.Class_Features("cuda", 13)
    ref = O.one_hot_with_overflow(torch.from_numpy(lab), 13, 13)
    assert torch.equal(cf.process_label(G(lab)).cpu(), ref)
    st = mods.dropin.SelfTrain(_args(), None, None, None, "cuda")
    ref = O.one_hot_with_overflow(torch.from_numpy(lab), 13, 12)
    assert torch.equal(st.process_label(G(lab)).cpu(), ref)
    with pytest.raises(mods.ops.TssError):
        mods.ops.one_hot_overflow(G(lab), 13, 255)                # scatter_ index out of range in the reference


def test_full2weak_and_prototype_weight(mods):
    r = rs(13)
    feat = r.standard_normal((2, 64, 40, 80)).astype(np.float32)
    protos = r.standard_normal((13, 64)).astype(np.float32)
    st = mods.dropin.SelfTrain(_args(), None, None, None, "cuda", dim=64)
    st.objective_vectors = G(protos)
    ref_f = O.full2weak(torch.from_numpy(feat), WEAK)
    out_f = st.full2weak(G(feat), WEAK)
    assert tuple(out_f.shape) == tuple(ref_f.shape) == (2, 64, 64, 128)
    close(out_f.cpu().numpy(), ref_f.numpy(), rtol=2e-6)
    ref_w = O.prototype_weight(torch.from_numpy(feat), torch.from_numpy(protos), 0.5, WEAK)
    ref_d = O.prototype_distance(ref_f, torch.from_numpy(protos))
    np.testing.assert_allclose(st.get_prototype_weight(G(feat), WEAK).cpu().numpy(), ref_w.numpy(),
                               rtol=weight_rtol(ref_d.numpy()), atol=1e-30)
    close(st.feat_prototype_distance(out_f).cpu().numpy(), ref_d.numpy())
    bad = {k: [t.clone() for t in v] for k, v in WEAK.items()}
    bad["RandomCrop"][1][1] += 8                                   # crops of different size: torch.cat raises
    with pytest.raises(RuntimeError):
        st.full2weak(G(feat), bad)


def test_selftrain_rectify_loss_and_prototype_update(mods):
    r = rs(14)
    n, c, k, h, w = 2, 13, 256, 16, 32
    a = _args(train_thred=0.3)
    st = mods.dropin.SelfTrain(a, None, None, None, "cuda")
    feat = r.standard_normal((n, k, h, w)).astype(np.float32)
    protos = r.standard_normal((c, k)).astype(np.float32)
    soft = torch.softmax(torch.from_numpy(3 * r.standard_normal((n, c, h, w)).astype(np.float32)), 1)
    hard = r.randint(0, c, size=(n, 1, h, w)).astype(np.int64)
    hard[r.rand(n, 1, h, w) < 0.6] = 12
    st.objective_vectors = torch.from_numpy(protos)               # the loader rebinds a CPU tensor (:72)
    w_ref = O.prototype_weight(torch.from_numpy(feat), torch.from_numpy(protos), 0.5)
    lab_ref, rect = O.rectify_labels(w_ref, soft, torch.from_numpy(hard), 0.3, 12)
    lab, wt = st.rectify_labels(G(feat), soft.cuda(), G(hard), want_weight=True)
    d_ref = O.prototype_distance(torch.from_numpy(feat), torch.from_numpy(protos)).numpy()
    np.testing.assert_allclose(wt.cpu().numpy(), w_ref.numpy(), rtol=weight_rtol(d_ref), atol=1e-30)
    top2 = rect.topk(2, dim=1)[0]
    eta = 2 * 0.5 * 2e-6 * float(d_ref.max())        # weight error propagated from distances within 2e-6 relative
    safe = ((top2[:, :1] - top2[:, 1:2]).abs() > 1e-6 + 2 * eta * top2[:, :1]) & \
           ((top2[:, :1] - 0.3).abs() > 1e-6 + eta * top2[:, :1])
    assert torch.equal(lab.cpu()[safe], lab_ref[safe]) and safe.float().mean() > 0.99
    # CE + RCE with gradient
    logits = torch.from_numpy(r.standard_normal((n, c, h, w)).astype(np.float32))
    tgt = lab_ref.reshape(n, h, w)
    x_ref = logits.clone().requires_grad_(True)
    ref = 0.1 * O.cross_entropy(x_ref, tgt, 12) + 1.0 * O.reverse_cross_entropy(x_ref, tgt, 13, 12)
    ref.backward()
    x = logits.cuda().requires_grad_(True)
    loss = st.target_loss(x, tgt.cuda())
    loss.backward()
    close(loss.item(), ref.item())
    close(x.grad.cpu().numpy(), x_ref.grad.numpy(), atol=1e-9)
    close(st.rce(logits.cuda(), tgt.cuda()).item(), O.reverse_cross_entropy(logits, tgt, 13, 12).item())
    close(st.seg_loss(logits.cuda(), tgt.cuda()).item(), O.cross_entropy(logits, tgt, 12).item())
    with pytest.raises(RuntimeError):
        st.rce(logits.cuda(), torch.full((n, h, w), 14, dtype=torch.int64).cuda())
    # moving-average prototype update of the batch (self_training.py:123-126)
    out = 3 * r.standard_normal((n, c, h, w)).astype(np.float32)
    st.update_prototypes(G(feat), G(out), start_mean=False)
    vecs, ids = O.class_mean_vectors(torch.from_numpy(feat), torch.from_numpy(out), 13, overflow_channel=12,
                                     use_conf_mask=True)
    bank = O.PrototypeBank(13, 256, momentum=1e-4)
    bank.vectors = torch.from_numpy(protos).clone()
    for t, v in zip(ids, vecs):
        bank.update(t, v, "moving_average", start_mean=False)
    close(st.objective_vectors.cpu().numpy(), bank.vectors.numpy())
    lv, li = st.calculate_mean_vector(G(feat), G(out))
    assert li == ids
    for v, rv in zip(lv, vecs):
        close(v.cpu().numpy(), rv.numpy())


# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("c", [13, 19])
def test_evaluate_matches_reference(mods, c):
    r = rs(20 + c)
    gts = [r.randint(0, c + 1, size=(64, 96)).astype(np.int64) for _ in range(5)]     # value c is masked out
    for g in gts:
        g[r.rand(*g.shape) < 0.02] = 255
    preds = [np.where(r.rand(*g.shape) < 0.7, np.minimum(g, c - 1), r.randint(0, c, size=g.shape)).astype(np.int64)
             for g in gts]
    ref = O.evaluate(preds, gts, c)
    for got in (mods.dropin.evaluate(preds, gts, c),                                  # list of per-image arrays
                mods.dropin.evaluate(np.stack(preds), np.stack(gts), c),              # what seg_validate passes
                mods.dropin.evaluate(G(np.stack(preds).astype(np.uint8)), G(np.stack(gts).astype(np.uint8)), c)):
        for x, y in zip(got[:4], ref[:4]):
            assert x == y                                                             # identical float64
        assert got[4].keys() == ref[4].keys() and all(got[4][k] == ref[4][k] or np.isnan(ref[4][k]) for k in ref[4])
    h = mods.dropin._fast_hist(preds[0].flatten(), gts[0].flatten(), c)
    assert h.dtype == np.int64 and np.array_equal(h, O.fast_hist(preds[0].flatten(), gts[0].flatten(), c))
    bad = preds[0].copy()
    bad[0, 0], gts[0][0, 0] = c + 3, 0                                                # np.bincount/reshape raises
    with pytest.raises(ValueError):
        mods.dropin._fast_hist(bad.flatten(), gts[0].flatten(), c)


def test_validate_batch_fused_argmax_histogram(mods):
    r = rs(31)
    logits = (3 * r.standard_normal((3, 13, 33, 65))).astype(np.float32)
    gt = r.randint(0, 14, size=(3, 256, 512)).astype(np.int64)
    meter = mods.dropin.ConfusionMeter(13)
    pred, loss = mods.dropin.validate_batch(G(logits), G(gt), meter, size=(256, 512), ignore_index=13)
    up = O.upsample_bilinear_ac(torch.from_numpy(logits), (256, 512))
    ref_pred = up.max(1)[1]
    top2 = up.topk(2, dim=1)[0]
    safe = (top2[:, 0] - top2[:, 1]) > 1e-6
    assert torch.equal(pred.cpu().long()[safe], ref_pred[safe]) and safe.float().mean() > 0.999
    close(loss.item(), O.cross_entropy(up, torch.from_numpy(gt), 13).item())
    if bool(safe.all()):
        assert np.array_equal(meter.result(), O.fast_hist(ref_pred.numpy().reshape(-1), gt.reshape(-1), 13))


def test_losses_modules_and_gradients(mods):
    r = rs(32)
    x0 = torch.from_numpy(r.standard_normal((2, 19, 12, 20)).astype(np.float32))
    t0 = torch.from_numpy(r.randint(0, 19, size=(2, 12, 20)).astype(np.int64))
    for reduction in ("mean", "sum"):
        xr = x0.clone().requires_grad_(True)
        tr = t0.clone()
        ref = O.semantic_consistency(xr, tr, (3, 7), reduction)
        ref.backward()
        x = x0.cuda().requires_grad_(True)
        t = t0.cuda()
        loss = mods.dropin.SemanticConsistency((3, 7), reduction)(x, t)
        loss.backward()
        close(loss.item(), ref.item())
        close(x.grad.cpu().numpy(), xr.grad.numpy(), atol=1e-9)
        assert torch.equal(t.cpu(), tr)                             # the in-place 255 rewrite is part of the contract
    alpha = [float(a) for a in r.uniform(0.2, 1.0, 19)]
    for gamma, al, avg in ((0, None, True), (2, None, True), (2, alpha, False)):
        xr = x0.clone().requires_grad_(True)
        ref = O.focal_loss(xr, t0, gamma, al, avg)
        ref.backward()
        x = x0.cuda().requires_grad_(True)
        loss = mods.dropin.FocalLoss(gamma, al, avg)(x, t0.cuda())
        loss.backward()
        close(loss.item(), ref.item())
        close(x.grad.cpu().numpy(), xr.grad.numpy(), atol=1e-9)


def test_upsample_and_its_adjoint(mods):
    r = rs(33)
    x0 = torch.from_numpy(r.standard_normal((2, 5, 9, 17)).astype(np.float32))
    g0 = torch.from_numpy(r.standard_normal((2, 5, 40, 70)).astype(np.float32))
    xr = x0.clone().requires_grad_(True)
    O.upsample_bilinear_ac(xr, (40, 70)).backward(g0)
    x = x0.cuda().requires_grad_(True)
    y = mods.dropin._Upsample.apply(x, (40, 70))
    y.backward(g0.cuda())
    close(y.detach().cpu().numpy(), O.upsample_bilinear_ac(x0, (40, 70)).numpy(), rtol=2e-6)
    close(x.grad.cpu().numpy(), xr.grad.numpy())


def test_sharded_prototypes_single_rank_equals_sequential(mods):
    r = rs(34)
    feat = r.standard_normal((4, 256, 11, 13)).astype(np.float32)
    logits = (3 * r.standard_normal((4, 13, 11, 13))).astype(np.float32)
    for mode, start_mean in ((mods.ops.UPDATE_MEAN, True), (mods.ops.UPDATE_MOVING_AVERAGE, False)):
        bank = mods.sharding.ShardedPrototypes(13, 256, "cuda", mode, start_mean, 0.05)
        protos = torch.zeros(13, 256, device="cuda")
        num = torch.zeros(13, device="cuda")
        for _ in range(3):
            means = mods.ops.class_means(G(feat), G(logits), 13)
            bank.update(means)
            mods.ops.prototype_update(protos, num, means, mode, start_mean, 0.05)
        close(bank.protos.cpu().numpy(), protos.cpu().numpy())
        assert torch.equal(bank.num, num)


# ------------------------------------------------------------------------------------------------
# BASELINE.json full sizes: size-independent properties (the oracle would take minutes there)
def test_full_size_properties_cfg2(mods):
    g = torch.Generator(device="cuda").manual_seed(5)
    n, c, h, w, H, W, k = 16, 13, 81, 241, 640, 1920, 256
    logits = 3 * torch.randn(n, c, h, w, generator=g, device="cuda")
    feat = torch.randn(n, k, h, w, generator=g, device="cuda")
    pl = mods.ops.pseudo_label(logits, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12, want_conf32=True)
    # (1) every image is independent: one image alone gives the same bytes
    one = mods.ops.pseudo_label(logits[7:8].contiguous(), size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12)
    assert torch.equal(one.label[0], pl.label[7]) and torch.equal(one.conf16[0], pl.conf16[7])
    assert torch.equal(one.label_thr[0], pl.label_thr[7])
    # (2) the threshold output is a function of (label, fp16 conf) alone: the stand-alone block agrees
    thr = mods.ops.threshold_labels(pl.label, pl.conf16, c, thr=0.3, keep=(5, 11), ignore_index=12)
    assert torch.equal(thr, pl.label_thr)
    assert torch.equal(pl.conf16, pl.conf32.to(torch.float16))
    # (3) at the source grid points the upsampled argmax is the low-resolution argmax
    ys = torch.arange(0, H, 71, device="cuda")[: 640 // 71 + 1]       # rows 0, 71, ...: 639*k/80 exact for k*8-ish
    low = mods.ops.pseudo_label(logits, mode=mods.ops.PL_UP_SOFTMAX)
    assert torch.equal(low.label[:, 0, 0], pl.label[:, 0, 0]) and torch.equal(low.label[:, -1, -1], pl.label[:, -1, -1])
    # (4) confusion matrix: total = number of counted pixels; matrix of (x, x) is diagonal
    gt = torch.randint(0, 14, (n, H, W), generator=g, device="cuda", dtype=torch.int16).to(torch.uint8)
    hist, bad = mods.ops.confusion_matrix(pl.label.reshape(-1), gt.reshape(-1), c)
    assert int(hist.sum()) == int((gt < c).sum()) and int(bad) == 0
    assert torch.equal(hist.sum(1), torch.bincount(gt.reshape(-1).long(), minlength=14)[:c])
    d, _ = mods.ops.confusion_matrix(pl.label.reshape(-1), pl.label.reshape(-1), c)
    assert torch.equal(d, torch.diag(torch.bincount(pl.label.reshape(-1).long(), minlength=c)))
    # (5) segmented sums: linear in feat, and the class sums add up to the plain sum over pixels
    m1 = mods.ops.class_means(feat, logits, c)
    m2 = mods.ops.class_means(2.0 * feat, logits, c)
    close(m2.sums.cpu().numpy(), 2.0 * m1.sums.cpu().numpy(), rtol=1e-6)
    close(m1.sums.sum(1).cpu().numpy(), feat.sum(dim=(2, 3)).cpu().numpy(), rtol=2e-5, atol=1e-3)
    assert int(m1.count_all.sum()) == n * h * w and torch.equal(m1.count_all, m1.count_gated)
    lab_low = logits.argmax(1).reshape(n, -1)
    assert torch.equal(m1.seg_label.long(), lab_low)


def test_update_from_vectors_only_equals_flagged_update(mods):
    """The multi-GPU exact exchange ships only the mean vectors (sharding.update_gathered): the update kernel
    must reach the same bank from vectors alone (invalid vectors are all-zero and the kernel's own
    `vector.sum() == 0` skip drops them) as from vectors + validity flags + sums."""
    r = rs(77)
    n, c, k, h, w = 6, 13, 256, 9, 17
    feat = r.randn(n, k, h, w).astype(np.float32)
    logits = (3 * r.randn(n, c, h, w)).astype(np.float32)
    m = mods.ops.class_means(G(feat), G(logits), c)
    for mode, start_mean in ((mods.ops.UPDATE_MEAN, True), (mods.ops.UPDATE_MOVING_AVERAGE, False)):
        p1, n1 = torch.zeros(c, k, device="cuda"), torch.zeros(c, device="cuda")
        p2, n2 = torch.zeros(c, k, device="cuda"), torch.zeros(c, device="cuda")
        mods.ops.prototype_update(p1, n1, m, mode, start_mean, 0.05)
        bare = mods.ops.ClassMeans(m.vectors, None, None, None, None, None, None)
        mods.ops.prototype_update(p2, n2, bare, mode, start_mean, 0.05)
        assert torch.equal(p1, p2) and torch.equal(n1, n2)
        assert float(n1.sum()) > 0
