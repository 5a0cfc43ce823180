"""GPU parity of the whole self-training step and of the pieces VERDICT r01 found untested:

* dropin.SelfTrain.step() on the network outputs the REFERENCE's own step() consumed (replay nets), against the
  golden losses, rectified labels, logit gradients and prototype bank (self_training.py:49-134);
* the 3000-cap of the running mean through tss_proto_update (cal_prototype.py:153-169);
* the rectify kernel against the golden labels, fed the features that produced the golden weights, in the
  materialised form (full2weak tensor) and in the fused sampled form (1/8-stride features in, labels out).

Tolerances: labels exact outside the margin rule (top-2 margin or threshold gap of the reference's own rectified
probabilities below the propagated weight error, see weight_rtol); losses / gradients / prototypes 1e-5 relative.
"""
import types

import numpy as np
import pytest
import torch

import inputs as gi
from oracle import label_path as O

pytestmark = pytest.mark.gpu
RTOL = 1e-5
TAU = 0.5


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import dropin, ops
    return types.SimpleNamespace(dropin=dropin, ops=ops)


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def G(a):
    return T(a).cuda()


def close_vec(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-30)
    err = np.abs(a - b).max()
    assert err <= rtol * scale, "max err %.3e vs scale %.3e (rel %.3e)" % (err, scale, err / scale)


def weight_rtol(d_ref, rtol_d=RTOL, tau=TAU):
    """Propagated bound for w = softmax(-tau (d - min d)): dw_c / w_c = -tau (dd_c - sum_j w_j dd_j), so
    |dw / w| <= 2 tau max|dd|; with distances within rtol_d relative, max|dd| <= rtol_d max(d).  """
    return 2.0 * tau * rtol_d * float(np.max(d_ref))


def check_rectified(lab, ref_lab, ref_weights, soft_up, train_thred, eta):
    """labels exact except where the reference's own rectified probabilities have a top-2 margin (argmax) or a
    threshold gap (ignore rule) inside the propagated weight error eta (relative) + 1e-6"""
    lab, ref_lab = np.asarray(lab).reshape(ref_lab.shape), np.asarray(ref_lab)
    bad = lab != ref_lab
    if not bad.any():
        return 0
    rect = ref_weights * soft_up
    rect = rect / rect.sum(1, keepdims=True)
    srt = np.sort(rect, axis=1)
    top, second = srt[:, -1], srt[:, -2]
    margin_ok = (top - second) < 1e-6 + 2 * eta * top
    gap_ok = np.abs(top - train_thred) < 1e-6 + eta * top
    assert (margin_ok | gap_ok)[bad].all(), "%d rectified labels differ outside the margin rule" % int(bad.sum())
    return int(bad.sum())


class _ReplayNet(torch.nn.Module):
    """Stands in for DeepLab: returns the recorded {'feat', 'out'} of successive forward passes (leaf tensors, so the
    step's backward() leaves the gradient on them)."""

    def __init__(self, passes):
        super().__init__()
        self.passes = passes
        self.i = 0

    def forward(self, x):
        d = self.passes[self.i % len(self.passes)]
        self.i += 1
        return dict(d)


def _args(**kw):
    base = dict(num_classes=13, ignore_index=12, proto_rectify=True, ema=True, rce=True, rce_alpha=0.1, rce_beta=1.0,
                proto_temperature=TAU, proto_momentum=1e-4, moving_prototype=True, train_thred=0)
    base.update(kw)
    return types.SimpleNamespace(**base)


def _weak():
    return {"RandomSized": [torch.tensor(v) for v in gi.WEAK_PARAMS["RandomSized"]],
            "RandomCrop": [torch.tensor(v) for v in gi.WEAK_PARAMS["RandomCrop"]],
            "RandomHorizontallyFlip": torch.tensor(gi.WEAK_PARAMS["RandomHorizontallyFlip"])}


@pytest.mark.parametrize("tag,train_thred", [("t0", 0), ("t06", 0.6)])
def test_selftrain_step_matches_reference_step(mods, golden, tag, train_thred):
    g = lambda k: golden["step_%s_%s" % (tag, k)]
    si = gi.step_inputs()
    src_out = G(g("net_src_out")).requires_grad_(True)
    tgt_out = G(g("net_tgt_out")).requires_grad_(True)
    feat_dummy = torch.zeros(2, 256, 8, 16, device="cuda")        # seg_net's feat is interpolated and never read (:76)
    seg = _ReplayNet([{"feat": feat_dummy, "out": src_out}, {"feat": feat_dummy, "out": tgt_out}])
    ema = _ReplayNet([{"feat": G(g("net_ema_feat")), "out": G(g("net_ema_out"))}])
    knob = torch.nn.Parameter(torch.zeros(1, device="cuda"))
    opt = torch.optim.SGD([knob], lr=0.0)
    st = mods.dropin.SelfTrain(_args(train_thred=train_thred), seg, ema, opt, "cuda")
    st.objective_vectors = T(gi.selftrain_protos()).clone()      # the loader rebinds a loaded tensor (:72)
    seen = {}
    real = st.target_loss
    st.target_loss = lambda logits, label: (seen.update(logits=logits.detach(), label=label.detach()), real(logits, label))[1]
    loss_t, loss_s = st.step(G(si["src_img"]), G(si["src_lab"]), G(si["tgt_img"]), G(si["tgt_full"]), G(si["lp_soft"]),
                             G(si["lp_hard"]), _weak(), None)
    torch.cuda.synchronize()
    np.testing.assert_allclose(loss_s.item(), g("loss_source"), rtol=RTOL)
    np.testing.assert_allclose(loss_t.item(), g("loss_target"), rtol=RTOL)
    close_vec(seen["logits"].cpu().numpy(), g("target_logits"), rtol=2e-6)
    # labels: the distances behind the golden weights are ~sqrt(2 K) = 22; the kernel's are within 1e-6 relative
    d_max = np.sqrt(2 * 256.0) * 1.3
    eta = 2 * TAU * 2e-6 * d_max
    flips = check_rectified(seen["label"].cpu().numpy(), g("rectified_label"), g("weights"), g("soft_up"), train_thred, eta)
    if flips == 0:
        close_vec(src_out.grad.cpu().numpy(), g("net_src_out_grad"))
        close_vec(tgt_out.grad.cpu().numpy(), g("net_tgt_out_grad"))
    else:   # a flipped label changes that pixel's gradient: compare the source branch only
        close_vec(src_out.grad.cpu().numpy(), g("net_src_out_grad"))
    # moving-average update (m = 1e-4): elementwise, as test_prototype_update_golden does for st_ma_vectors
    np.testing.assert_allclose(st.objective_vectors.cpu().numpy(), g("protos_after"), rtol=1e-6, atol=1e-7)
    assert np.array_equal(st.objective_vectors_num.cpu().numpy(), g("num_after"))
    assert st.counter == 1


@pytest.mark.parametrize("tag,train_thred", [("t0", 0.0), ("t06", 0.6)])
@pytest.mark.parametrize("path", ["materialised", "sampled"])
def test_rectify_kernels_reproduce_golden_labels(mods, golden, tag, train_thred, path):
    """The features behind the golden step weights are the EMA net's 1/8-stride map interpolated to the quarter
    resolution (self_training.py:69); feed the distance / rectify kernels exactly those."""
    ops = mods.ops
    g = lambda k: golden["step_%s_%s" % (tag, k)]
    si = gi.step_inputs()
    protos = G(gi.selftrain_protos())
    soft_up = G(g("soft_up"))
    hard = ops.resize_nearest_u8(G(si["lp_hard"].astype(np.uint8)), (16, 32), rule="torch")
    assert np.array_equal(hard.cpu().numpy().astype(np.int64), O.resize_nearest_torch(T(si["lp_hard"]), (16, 32)).numpy())
    ema_feat8 = T(g("net_ema_feat"))
    if path == "materialised":
        w_in = O.upsample_bilinear_ac(ema_feat8, (20, 40))
        feat_weak = O.full2weak(w_in, _weak()).cuda()
        d, wgt, lab = ops.prototype_distance(feat_weak, protos, temperature=TAU, want_dist=True, want_weight=True, soft=soft_up,
                                             hard=hard, train_thred=train_thred, ignore_index=12)
        d_ref = O.prototype_distance(feat_weak.cpu(), protos.cpu()).numpy()
    else:
        d, wgt, lab = ops.prototype_distance_sampled(ema_feat8.cuda(), (20, 40), _weak(), protos, temperature=TAU, want_dist=True,
                                                     want_weight=True, soft=soft_up, hard=hard, train_thred=train_thred,
                                                     ignore_index=12)
        d_ref = O.prototype_distance(O.full2weak(O.upsample_bilinear_ac(ema_feat8, (20, 40)), _weak()), protos.cpu()).numpy()
    np.testing.assert_allclose(d.cpu().numpy(), d_ref, rtol=RTOL)
    rel_d = float(np.max(np.abs(d.cpu().numpy() - d_ref) / d_ref))
    np.testing.assert_allclose(wgt.cpu().numpy(), g("weights"), rtol=weight_rtol(d_ref), atol=1e-30)
    eta = 2 * TAU * max(rel_d, 1e-7) * float(d_ref.max())
    check_rectified(lab.cpu().numpy(), g("rectified_label"), g("weights"), g("soft_up"), train_thred, eta)


def test_running_mean_cap_3000(mods, golden):
    """3100 sequential 'mean' updates of one class: num saturates at 3000 (cal_prototype.py:165-168) and the vector
    follows the reference's fp32 operation sequence."""
    ops = mods.ops
    many = gi.cap_vectors()                                         # the vectors make_golden.py fed the reference
    protos = torch.zeros(13, 256, device="cuda")
    num = torch.zeros(13, device="cuda")
    # one launch: [3100, 1, 256] vectors into row 6 (the single-class form update_objective_SingleVector uses)
    row, cnt = protos[6:7], num[6:7]
    means = ops.ClassMeans(G(many).reshape(-1, 1, 256), None, None, None, None, None, None)
    ops.prototype_update(row, cnt, means, ops.UPDATE_MEAN, True, 0.9999)
    assert np.array_equal(num.cpu().numpy(), golden["cf_cap_num"]) and float(num[6]) == 3000.0
    close_vec(protos[6].cpu().numpy(), golden["cf_cap_vector"], rtol=1e-6)
    # the same through 3100 single-vector calls of the drop-in method
    cf = mods.dropin.Class_Features("cuda", 13)
    dev_many = G(many)
    for j in range(many.shape[0]):
        cf.update_objective_SingleVector(6, dev_many[j].reshape(256, 1, 1), "mean")
    assert np.array_equal(cf.objective_vectors_num.cpu().numpy(), golden["cf_cap_num"])
    assert torch.equal(cf.objective_vectors[6], protos[6])


def test_rce_accepts_label_equal_to_num_classes(mods):
    """one_hot(label, C+1)[:, :C] (self_training.py:171-172) accepts label == C: the pixel joins the mask and adds
    -log(1e-4) sum_c clamp(p_c); larger labels raise.  """
    r = gi.rs(2100)
    logits = T((2 * r.standard_normal((2, 13, 8, 16))).astype(np.float32))
    label = T(r.randint(0, 14, size=(2, 8, 16)).astype(np.int64))    # 13 == C appears, 12 == ignore appears
    assert int((label == 13).sum()) > 0
    ref = O.reverse_cross_entropy(logits.clone().requires_grad_(True), label.clone(), 13, 12)
    st = mods.dropin.SelfTrain(_args(), None, None, None, "cuda")
    x = logits.cuda().requires_grad_(True)
    got = st.rce(x, label.cuda())
    np.testing.assert_allclose(got.item(), ref.item(), rtol=RTOL)
    xr = logits.clone().requires_grad_(True)
    O.reverse_cross_entropy(xr, label.clone(), 13, 12).backward()
    got.backward()
    close_vec(x.grad.cpu().numpy(), xr.grad.numpy())
    with pytest.raises(RuntimeError):
        st.rce(logits.cuda(), torch.full((2, 8, 16), 14, dtype=torch.int64).cuda())


def test_strict_label_mode_raises_like_torch(mods, monkeypatch):
    logits = G(gi.randn(2101, 1, 13, 4, 4))
    tgt = torch.full((1, 4, 4), 3, dtype=torch.int64, device="cuda")
    tgt[0, 0, 0] = 40
    loss = mods.dropin._SegLoss(12)
    assert torch.isfinite(loss(logits, tgt))                        # default: counted, not raised (no host sync)
    monkeypatch.setattr(mods.dropin, "STRICT_LABELS", True)
    with pytest.raises(IndexError):
        loss(logits, tgt)
