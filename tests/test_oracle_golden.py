"""Pins the CPU oracle (oracle/label_path.py) to the reference: every array in
tests/golden/reference_golden.npz is an output of the UNMODIFIED reference run on CPU
(tests/golden/make_golden.py); inputs are regenerated from seeds (tests/golden/inputs.py)."""
import numpy as np
import pytest
import torch

import inputs as gi
from oracle import label_path as O

torch.set_num_threads(4)


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.parametrize("name", list(gi.PL_CASES))
def test_pseudo_label_branches(golden, name):
    n, c, h, w, H, W, _ = gi.PL_CASES[name]
    logits, flip_logits = gi.pl_logits(name)
    assert np.array_equal(O.pseudo_label_soft(T(logits)).numpy(), golden[name + "_soft_prob"])
    conf, lab = O.pseudo_label_hard(T(logits))
    assert np.array_equal(lab.numpy().astype(np.uint8), golden[name + "_hard_label"])
    assert np.array_equal(conf.numpy().astype(np.float16), golden[name + "_hard_conf"])
    conf, lab = O.pseudo_label_flip(T(logits), T(flip_logits), (H, W))
    assert np.array_equal(lab.numpy().astype(np.uint8), golden[name + "_flip_label"])
    assert np.array_equal(conf.numpy().astype(np.float16), golden[name + "_flip_conf"])
    assert np.array_equal(O.eval_prediction(T(logits), (H, W)).numpy(), golden[name + "_eval_pred"])
    up = O.upsample_bilinear_ac(T(logits), (H, W)).numpy()
    assert np.array_equal(up[:, :3], golden[name + "_up_logits_c0to2"])
    # explicit-formula restatement (what the kernel implements): tight, not bit-exact
    up_np = O.upsample_bilinear_ac_np(logits, (H, W))
    assert np.abs(up_np - up).max() <= 2e-6
    assert np.array_equal(up_np.argmax(1), golden[name + "_eval_pred"])


def test_threshold_block(golden):
    label, conf = gi.threshold_inputs()
    kept = O.threshold_hard_label(label, conf.astype(np.float16), 0.3, 12, keep=(5, 11))
    out = O.resize_nearest_pil_np(kept, (700, 320))
    assert np.array_equal(out, golden["thr_label_hard_700x320"])
    # per-class vector form with the reference's scalar as a special case
    kept2 = O.threshold_hard_label(label, conf.astype(np.float16), np.full(13, 0.3), 12, keep=(5, 11))
    assert np.array_equal(kept, kept2)
    # the fp16 compare really matters on the planted band
    naive = label.copy()
    naive[conf <= np.float32(0.3)] = 12
    assert (naive[0, :13] != np.where(kept[0, :13] == 12, 12, label[0, :13])).any()


def test_class_features(golden):
    feat, logits, labels_val = gi.proto_inputs()
    vecs, ids = O.class_mean_vectors(T(feat), T(logits), 13)
    assert np.array_equal(np.array(ids), golden["cf_ids"])
    assert np.array_equal(np.stack([v.reshape(-1).numpy() for v in vecs]), golden["cf_vectors"])
    bank = O.PrototypeBank(13)
    for i, v in zip(ids, vecs):
        bank.update(i, v, "mean")
    feat2 = gi.randn(302, *feat.shape)
    vecs2, ids2 = O.class_mean_vectors(T(feat2), T(logits), 13)
    for i, v in zip(ids2, vecs2):
        bank.update(i, v, "mean")
    assert np.array_equal(bank.vectors.numpy(), golden["cf_objective_vectors"])
    assert np.array_equal(bank.num.numpy(), golden["cf_objective_num"])
    vecs3, ids3 = O.class_mean_vectors(T(feat), T(logits), 13, labels_val=T(labels_val))
    assert np.array_equal(np.array(ids3), golden["cf_val_ids"])
    assert np.array_equal(np.stack([v.reshape(-1).numpy() for v in vecs3]), golden["cf_val_vectors"])
    hot = O.one_hot_with_overflow(T(labels_val), 13, 13).numpy().astype(np.uint8)
    assert np.array_equal(hot, golden["cf_process_label"])


def test_prototype_cap_and_closed_form(golden):
    many = gi.cap_vectors()
    bank = O.PrototypeBank(13)
    for j in range(many.shape[0]):
        bank.update(6, T(many[j]).reshape(256, 1, 1), "mean")
    assert np.array_equal(bank.vectors[6].numpy(), golden["cf_cap_vector"])
    assert np.array_equal(bank.num.numpy(), golden["cf_cap_num"])
    vec, num = O.closed_form_update(np.zeros(256), 0.0, many, "mean", 0.0)
    assert num == 3000.0
    scale = np.abs(golden["cf_cap_vector"]).max()
    assert np.abs(vec - golden["cf_cap_vector"]).max() <= 1e-5 * scale


def _st_bank():
    bank = O.PrototypeBank(13, momentum=1e-4)
    bank.vectors = T(gi.selftrain_protos()).clone()
    return bank


def test_selftrain_methods(golden):
    protos = T(gi.selftrain_protos())
    feat = T(gi.selftrain_feat())
    assert np.array_equal(O.prototype_distance(feat, protos).numpy(), golden["st_distance"])
    ff = T(gi.full2weak_feat())
    f2w = O.full2weak(ff, gi.WEAK_PARAMS)
    assert np.array_equal(f2w.numpy()[:, ::16], golden["st_full2weak_k16"])
    assert np.array_equal(O.prototype_weight(ff, protos, 0.5, gi.WEAK_PARAMS).numpy(), golden["st_weight"])
    logits, label = gi.ce_inputs()
    lg = T(logits).requires_grad_(True)
    ce = O.cross_entropy(lg, T(label), 12)
    rce = O.reverse_cross_entropy(lg, T(label).clone(), 13, 12)
    (0.1 * ce + 1.0 * rce).backward()
    assert np.array_equal(ce.detach().numpy(), golden["st_ce"])
    assert np.array_equal(rce.detach().numpy(), golden["st_rce"])
    assert np.array_equal(lg.grad.numpy(), golden["st_loss_grad"])
    pfeat, plogits, _ = gi.proto_inputs()
    for tag, th in (("none", None), ("t05", 0.5)):
        vecs, ids = O.class_mean_vectors(T(pfeat), T(plogits), 13, thresh=th, overflow_channel=12, use_conf_mask=True)
        assert np.array_equal(np.array(ids), golden["st_mv_%s_ids" % tag])
        assert np.array_equal(np.stack([v.reshape(-1).numpy() for v in vecs]), golden["st_mv_%s_vectors" % tag])
    vecs, ids = O.class_mean_vectors(T(pfeat), T(plogits), 13, overflow_channel=12, use_conf_mask=True)
    bank = _st_bank()
    for i, v in zip(ids, vecs):
        bank.update(i, v, start_mean=False)
    assert np.array_equal(bank.vectors.numpy(), golden["st_ma_vectors"])
    assert np.array_equal(bank.num.numpy(), golden["st_ma_num"])
    bank = O.PrototypeBank(13, momentum=1e-4)
    for i, v in zip(ids, vecs):
        bank.update(i, v)
    assert np.array_equal(bank.vectors.numpy(), golden["st_sm_vectors"])
    assert np.array_equal(bank.num.numpy(), golden["st_sm_num"])


@pytest.mark.parametrize("tag,train_thred", [("t0", 0), ("t06", 0.6)])
def test_step_rectify_and_loss(golden, tag, train_thred):
    """The rectify block (self_training.py:103-112) and the CE+RCE target loss (:114-118), fed with the
    tensors the reference's own step() produced (weights, upsampled soft labels, target logits)."""
    si = gi.step_inputs()
    weights = T(golden["step_%s_weights" % tag])
    soft_up = O.upsample_bilinear_ac(T(si["lp_soft"]), (16, 32))
    assert np.array_equal(soft_up.numpy(), golden["step_%s_soft_up" % tag])
    hard = O.resize_nearest_torch(T(si["lp_hard"]), (16, 32))
    lab, _ = O.rectify_labels(weights, soft_up, hard, train_thred, 12)
    assert np.array_equal(lab[:, 0].numpy(), golden["step_%s_rectified_label" % tag])
    lg = T(golden["step_%s_target_logits" % tag])
    tgt = lab[:, 0]
    loss = 0.1 * O.cross_entropy(lg, tgt, 12) + 1.0 * O.reverse_cross_entropy(lg, tgt.clone(), 13, 12)
    assert np.array_equal(loss.numpy(), golden["step_%s_loss_target" % tag])


def _weak_params():
    return {"RandomSized": [torch.tensor(v) for v in gi.WEAK_PARAMS["RandomSized"]],
            "RandomCrop": [torch.tensor(v) for v in gi.WEAK_PARAMS["RandomCrop"]],
            "RandomHorizontallyFlip": torch.tensor(gi.WEAK_PARAMS["RandomHorizontallyFlip"])}


@pytest.mark.parametrize("tag,train_thred", [("t0", 0), ("t06", 0.6)])
def test_whole_step_from_recorded_network_outputs(golden, tag, train_thred):
    """SelfTrain.step() (self_training.py:49-134) restated on the network outputs the reference's step consumed:
    both losses, the rectified labels, the gradients its two backward() calls left on the raw logits, and the
    prototype bank after the moving-average update."""
    si = gi.step_inputs()
    g = lambda k: T(golden["step_%s_%s" % (tag, k)])
    out = O.self_train_step(g("net_src_out"), g("net_ema_feat"), g("net_ema_out"), g("net_tgt_out"), T(si["src_lab"]),
                            si["tgt_full"].shape[2:], T(si["lp_soft"]), T(si["lp_hard"]), _weak_params(),
                            T(gi.selftrain_protos()), torch.zeros(13), train_thred=train_thred)
    for k in ("loss_target", "loss_source", "rectified_label", "weights", "soft_up", "target_logits", "protos_after",
              "num_after"):
        assert np.array_equal(out[k].numpy(), golden["step_%s_%s" % (tag, k)]), k
    assert np.array_equal(out["src_out_grad"].numpy(), golden["step_%s_net_src_out_grad" % tag])
    assert np.array_equal(out["tgt_out_grad"].numpy(), golden["step_%s_net_tgt_out_grad" % tag])


@pytest.mark.parametrize("c", [13, 19])
def test_eval_tools(golden, c):
    pred, gt = gi.hist_inputs(seed=601 + c, c=c)
    assert np.array_equal(O.fast_hist(pred.flatten(), gt.flatten(), c), golden["hist_c%d" % c])
    acc, acc_cls, miou, fwav, cls_iu = O.evaluate(list(pred), list(gt), c)
    assert np.array_equal(np.array([acc, acc_cls, miou, fwav]), golden["eval_c%d" % c])
    assert np.array_equal(np.array(list(cls_iu.values())), golden["eval_c%d_cls_iu" % c], equal_nan=True)
    with pytest.raises(ValueError):
        O.evaluate(list(np.zeros((1, 4), np.int64)), list(np.zeros((1, 4), np.int64)), 7)


def test_losses(golden):
    import warnings
    logits, label = gi.ce_inputs(seed=502, c=13)
    lg = T(logits).requires_grad_(True)
    tgt = T(label).clone()
    loss = O.semantic_consistency(lg, tgt, (3, 5))
    loss.backward()
    assert np.array_equal(loss.detach().numpy(), golden["sc_loss"])
    assert np.array_equal(lg.grad.numpy(), golden["sc_grad"])
    assert np.array_equal(tgt.numpy(), golden["sc_target_after"])
    for tag, kw in (("g2", dict(gamma=2)), ("g2a", dict(gamma=2, alpha=[0.5 + 0.05 * i for i in range(13)])),
                    ("g0sum", dict(gamma=0, size_average=False))):
        lg = T(logits).requires_grad_(True)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            fl = O.focal_loss(lg, T(label), **kw)
        fl.backward()
        assert np.array_equal(fl.detach().numpy(), golden["focal_%s_loss" % tag])
        assert np.array_equal(lg.grad.numpy(), golden["focal_%s_grad" % tag])
