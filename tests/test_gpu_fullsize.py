"""Oracle comparison at the REAL sizes of the BASELINE.json configurations: the kernels run the
whole configuration, the oracle (the reference's own torch / numpy op sequence) checks one or two images of it --
seconds of CPU each -- so that grid sizing, tile edges and row bookkeeping are pinned at the benchmarked shapes
and not only at the small golden shapes.

cfg2: 16 x 13 x 81 x 241 -> 640 x 1920 (pseudo-labels, 256-ch prototypes, confusion matrix, CE + RCE)
cfg3: 129 x 257 features, K = 2048, C = 19 (class means, distances, weights, rectified labels)
cfg5: 256 x 512 uint8 label maps, C = 19 (confusion matrix + metrics)
"""
import numpy as np
import pytest
import torch

from oracle import label_path as O

pytestmark = pytest.mark.gpu
RTOL = 1e-5


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import ops as _ops
    return _ops


def close_vec(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max()
    assert err <= rtol * scale, "max err %.3e vs scale %.3e (rel %.3e)" % (err, scale, err / scale)


def _cfg2_inputs():
    g = torch.Generator().manual_seed(20260)
    logits = 3.0 * torch.randn(16, 13, 81, 241, generator=g)
    feat = torch.randn(16, 256, 81, 241, generator=g)
    gt = torch.randint(0, 13, (16, 640, 1920), generator=g, dtype=torch.int64).to(torch.uint8)
    return logits, feat, gt


def test_cfg2_full_size_against_oracle(ops):
    logits, feat, gt = _cfg2_inputs()
    H, W, c = 640, 1920, 13
    dl, df, dg = logits.cuda(), feat.cuda(), gt.cuda()
    pl = ops.pseudo_label(dl, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12, want_conf32=True)
    # the co-resident form of the same kernel (12 warps per SM) must give the same bytes
    pl12 = ops.pseudo_label(dl, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12, flags=ops.pl_flags(sm_warps=12))
    assert torch.equal(pl.label, pl12.label) and torch.equal(pl.conf16, pl12.conf16) and torch.equal(pl.label_thr, pl12.label_thr)
    # ... and so must the multi-wave small-CTA form, the two-rows-per-instruction-stream form and the 96-register build
    for fl in (ops.CORESIDENT_PL_FLAGS, ops.pl_flags(rows2=True), ops.pl_flags(sm_warps=12, cta_warps=3, waves=2, rows2=True),
               ops.pl_flags(regs96=True), ops.pl_flags(sm_warps=20, cta_warps=4, waves=2, regs96=True)):
        alt = ops.pseudo_label(dl, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12, flags=fl)
        assert torch.equal(pl.label, alt.label) and torch.equal(pl.conf16, alt.conf16) and torch.equal(pl.label_thr, alt.label_thr)
    thr16 = float(np.float16(0.3))
    for i in (0, 9, 15):                                            # first, an interior and the last image of the batch
        up = O.upsample_bilinear_ac(logits[i:i + 1], (H, W))
        prob = torch.softmax(up, 1)
        conf_ref, lab_ref = prob.max(1)
        top2 = prob.topk(2, dim=1)[0]
        safe = ((top2[:, 0] - top2[:, 1]) > 1e-6)[0].numpy()
        lab = pl.label[i].cpu().numpy()
        assert (lab[safe] == lab_ref[0].numpy()[safe]).all() and safe.mean() > 0.9999
        same = lab == lab_ref[0].numpy()
        c32 = pl.conf32[i].cpu().numpy()
        assert np.abs(c32 - conf_ref[0].numpy())[same].max() <= 2e-6
        # thresholded labels: the reference's block on ITS stored (uint8 label, fp16 conf); pixels whose confidence is
        # within 1e-6 of the fp16 rounding boundary of the comparison (or whose label differed) are exempt
        c16_ref = conf_ref[0].numpy().astype(np.float16)
        thr_ref = O.threshold_hard_label(lab_ref[0].numpy().astype(np.uint8), c16_ref, 0.3, 12, keep=(5, 11))
        boundary = thr16 + 2.0 ** -13                                # midpoint between fp16(0.3) and the next fp16 value
        exempt = ~same | (np.abs(conf_ref[0].numpy() - boundary) < 1e-6)
        got_thr = pl.label_thr[i].cpu().numpy()
        assert (got_thr[~exempt] == thr_ref[~exempt]).all()
        # fp16 confidence as stored: at most one fp16 ulp away, on well under 1 % of the pixels
        ulp = np.abs(pl.conf16[i].cpu().numpy().view(np.int16).astype(np.int32) - c16_ref.view(np.int16).astype(np.int32))
        assert ulp.max() <= 1 and (ulp > 0).mean() < 0.005
    # prototypes: class means of two images against the reference loop, then the update against the oracle bank
    m = ops.class_means(df, dl, c)
    for i in (0, 15):
        vr, ir = O.class_mean_vectors(feat[i:i + 1], logits[i:i + 1], c)
        valid = m.valid[i].cpu().numpy().astype(bool)
        assert list(np.nonzero(valid)[0]) == ir
        for t, v in zip(ir, vr):
            close_vec(m.vectors[i, t].cpu().numpy(), v.reshape(-1).numpy())
    # the 4-consumer (co-resident) form of the accumulate kernel gives the same sums up to fp32 summation order
    m4 = ops.class_means(df, dl, c, consumer_warps=4)
    close_vec(m4.sums.cpu().numpy(), m.sums.cpu().numpy(), rtol=2e-6)
    assert torch.equal(m4.count_all, m.count_all) and torch.equal(m4.valid, m.valid)
    # confusion matrix of two images against np.bincount
    for i in (3, 12):
        h, bad = ops.confusion_matrix(pl.label[i].reshape(-1), dg[i].reshape(-1), c)
        ref = O.fast_hist(pl.label[i].cpu().numpy().reshape(-1).astype(np.int64), gt[i].numpy().reshape(-1).astype(np.int64), c)
        assert int(bad) == 0 and np.array_equal(h.cpu().numpy(), ref)
    # CE + RCE with gradient on one full-resolution frame (13 x 640 x 1920), int64 targets as the reference passes
    g = torch.Generator().manual_seed(20261)
    x0 = torch.randn(1, c, H, W, generator=g)
    y0 = torch.randint(0, c, (1, H, W), generator=g)
    xr = x0.clone().requires_grad_(True)
    ref = 0.1 * O.cross_entropy(xr, y0, 12) + 1.0 * O.reverse_cross_entropy(xr, y0.clone(), c, 12)
    ref.backward()
    x = x0.cuda().requires_grad_(True)
    tot = ops.ce_rce(x, y0.cuda(), 12, 0.1, 1.0)[0]
    tot.backward()
    np.testing.assert_allclose(tot.item(), ref.item(), rtol=RTOL)
    close_vec(x.grad.cpu().numpy(), xr.grad.numpy())


def test_cfg3_full_size_against_oracle(ops):
    n, k, c, h, w = 2, 2048, 19, 129, 257
    g = torch.Generator().manual_seed(20263)
    feat = torch.randn(n, k, h, w, generator=g)
    logits = 3.0 * torch.randn(n, c, h, w, generator=g)
    protos = torch.randn(c, k, generator=g)
    soft = torch.softmax(3.0 * torch.randn(n, c, h, w, generator=g), 1)
    hard = torch.full((n, 1, h, w), 255, dtype=torch.int64)
    hard[:, :, ::7, ::5] = 4
    df = feat.cuda()
    m = ops.class_means(df, logits.cuda(), c)
    vr, ir = O.class_mean_vectors(feat[1:2], logits[1:2], c)                      # one image: ~6 s of CPU
    assert list(np.nonzero(m.valid[1].cpu().numpy())[0]) == ir
    for t, v in zip(ir, vr):
        close_vec(m.vectors[1, t].cpu().numpy(), v.reshape(-1).numpy())
    d_ref = O.prototype_distance(feat[1:2], protos).numpy()                        # ~10 s of CPU
    w_ref = O.prototype_weight(feat[1:2], protos, 0.5)
    lab_ref, rect = O.rectify_labels(w_ref, soft[1:2], hard[1:2], 0.3, 255)
    for form in (0, ops.dist_form(algo=ops.DIST_DIRECT)):
        d, wgt, lab = ops.prototype_distance(df, protos.cuda(), temperature=0.5, want_dist=True, want_weight=True,
                                             soft=soft.cuda(), hard=hard.cuda(), train_thred=0.3, ignore_index=255, form=form)
        got = d[1:2].cpu().numpy()
        np.testing.assert_allclose(got, d_ref, rtol=RTOL)
        rel_d = float(np.max(np.abs(got - d_ref) / d_ref))
        np.testing.assert_allclose(wgt[1:2].cpu().numpy(), w_ref.numpy(), rtol=2 * 0.5 * RTOL * float(d_ref.max()), atol=1e-30)
        bad = (lab[1:2].cpu() != lab_ref).numpy()
        if bad.any():
            r = rect.numpy()
            srt = np.sort(r, axis=1)
            top, second = srt[:, -1:], srt[:, -2:-1]
            eta = 2 * 0.5 * max(rel_d, 1e-7) * float(d_ref.max())
            ok = ((top - second) < 1e-6 + 2 * eta * top) | (np.abs(top - 0.3) < 1e-6 + eta * top)
            assert ok[bad].all() and bad.mean() < 1e-3


def test_cfg5_full_size_against_oracle(ops):
    from b200seg import dropin
    c, m = 19, 64
    r = np.random.RandomState(20265)
    gt = r.randint(0, 20, size=(m, 256, 512)).astype(np.uint8)                     # 19 = masked out
    gt[r.uniform(size=gt.shape) < 0.01] = 255
    pred = np.where(r.uniform(size=gt.shape) < 0.7, np.minimum(gt, c - 1), r.randint(0, c, size=gt.shape)).astype(np.uint8)
    ref = O.evaluate(list(pred.astype(np.int64)), list(np.where(gt == 255, -1, gt).astype(np.int64)), c)
    got = dropin.evaluate(torch.from_numpy(pred).cuda(), torch.from_numpy(gt).cuda(), c)
    for x, y in zip(got[:4], ref[:4]):
        assert x == y                                                              # identical float64
    # what seg_validate really passes: int64 predictions [M,H,W] and labels with a channel axis [M,1,H,W]
    got2 = dropin.evaluate(pred[:8].astype(np.int64), gt[:8].astype(np.int64)[:, None], c)
    ref2 = O.evaluate(list(pred[:8].astype(np.int64)), list(np.where(gt[:8] == 255, -1, gt[:8]).astype(np.int64)), c)
    for x, y in zip(got2[:4], ref2[:4]):
        assert x == y
    h = ops.confusion_matrix(torch.from_numpy(pred).cuda().reshape(-1), torch.from_numpy(gt).cuda().reshape(-1), c)[0]
    assert np.array_equal(h.cpu().numpy(), O.fast_hist(pred.reshape(-1).astype(np.int64),
                                                        np.where(gt == 255, -1, gt).reshape(-1).astype(np.int64), c))


@pytest.mark.parametrize("shape", [(16, 13, 81, 241, 640, 1920), (2, 19, 129, 257, 1024, 2048), (3, 13, 41, 81, 320, 640),
                                   (2, 13, 40, 80, 321, 642), (2, 7, 41, 81, 320, 640)])
def test_pseudo_label_with_fused_confusion_matrix(ops, shape):
    """tss_pseudo_label_hist: same labels / confidences as tss_pseudo_label, and the matrix np.bincount gives on them
    (utils/eval_tools.py:4-9), for the register-resident kernel (first three shapes, every launch form) and for the shapes
    that fall back to the histogram kernel (odd upsampling factor with 13 classes is still fused; 7 classes is not)."""
    n, c, h, w, H, W = shape
    g = torch.Generator().manual_seed(sum(shape))
    logits = (3.0 * torch.randn(n, c, h, w, generator=g)).cuda()
    gt = torch.randint(0, c + 2, (n, H, W), generator=g).to(torch.uint8)       # c, c + 1 = not a class: left out
    gt[:, ::5, ::3] = 255
    dg = gt.cuda()
    base = ops.pseudo_label(logits, size=(H, W), thr=0.3, ignore_index=c - 1)
    ref = O.fast_hist(base.label.cpu().numpy().reshape(-1).astype(np.int64),
                      np.where(gt.numpy() >= c, -1, gt.numpy()).reshape(-1).astype(np.int64), c)
    for fl in (0, ops.CORESIDENT_PL_FLAGS, ops.pl_flags(sm_warps=12, cta_warps=3, waves=3)):
        hist = torch.zeros(c, c, dtype=torch.int64, device="cuda")
        got = ops.pseudo_label(logits, size=(H, W), thr=0.3, ignore_index=c - 1, flags=fl, gt=dg, hist=hist)
        assert torch.equal(got.label, base.label) and torch.equal(got.conf16, base.conf16) and torch.equal(got.label_thr, base.label_thr)
        assert np.array_equal(hist.cpu().numpy(), ref)
        # accumulated, not overwritten
        ops.pseudo_label(logits, size=(H, W), thr=0.3, ignore_index=c - 1, flags=fl, gt=dg, hist=hist)
        assert np.array_equal(hist.cpu().numpy(), 2 * ref)
    # argmax-only mode (segmentation_evaluate.py:39,42) without the confidence output
    hist = torch.zeros(c, c, dtype=torch.int64, device="cuda")
    got = ops.pseudo_label(logits, size=(H, W), mode=ops.PL_UP_ARGMAX, want_conf16=False, gt=dg, hist=hist)
    assert torch.equal(got.label, base.label) and np.array_equal(hist.cpu().numpy(), ref)
