"""Live cross-check of the oracle against the UNMODIFIED reference, on inputs the golden fixtures do not contain
(fresh seeds, other shapes).  Runs wherever the reference tree is importable: /root/reference in the dev container,
baseline/_ref/ on the GPU box (staged by __graft_entry__.build()); skipped otherwise.  CPU only."""
import types

import numpy as np
import pytest
import torch

import ref_import
from oracle import label_path as O

pytestmark = pytest.mark.skipif(not ref_import.reference_available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref():
    return ref_import.import_reference()


def _args(**kw):
    base = dict(num_classes=13, ignore_index=12, proto_rectify=True, ema=True, rce=True, rce_alpha=0.1, rce_beta=1.0,
                proto_temperature=0.5, proto_momentum=1e-4, moving_prototype=True, train_thred=0)
    base.update(kw)
    return types.SimpleNamespace(**base)


def test_class_features_and_updates_live(ref):
    g = torch.Generator().manual_seed(31)
    feat = torch.randn(2, 256, 11, 19, generator=g)
    logits = 3 * torch.randn(2, 13, 11, 19, generator=g)
    cf = ref.cal_prototype.Class_Features(torch.device("cpu"), numbers=13)
    vecs, ids = cf.calculate_mean_vector(feat, logits)
    ovecs, oids = O.class_mean_vectors(feat, logits, 13)
    assert ids == oids and all(torch.equal(a, b) for a, b in zip(vecs, ovecs))
    bank = O.PrototypeBank(13, 256)
    for t, v in zip(ids, vecs):
        cf.update_objective_SingleVector(t, v, "mean")
        bank.update(t, v, "mean")
    assert torch.equal(cf.objective_vectors, bank.vectors) and torch.equal(cf.objective_vectors_num, bank.num)
    # closed form of the same recurrence (what the sharded update kernel evaluates): within fp32 rounding
    for c in set(ids):
        contribs = np.stack([v.reshape(-1).numpy() for t, v in zip(ids, vecs) if t == c])
        vec, n = O.closed_form_update(np.zeros(256), 0.0, contribs, "mean", 0.9999)
        assert n == bank.num[c].item()
        assert np.abs(vec - bank.vectors[c].numpy()).max() <= 1e-5 * np.abs(vec).max()


def test_selftrain_methods_live(ref):
    g = torch.Generator().manual_seed(32)
    st = ref.self_training.SelfTrain(_args(), None, None, None, torch.device("cpu"), None)
    protos = torch.randn(13, 256, generator=g)
    st.objective_vectors = protos.clone()
    feat = torch.randn(2, 256, 9, 13, generator=g)
    assert torch.equal(st.feat_prototype_distance(feat), O.prototype_distance(feat, protos))
    logits = 2 * torch.randn(2, 13, 9, 13, generator=g)
    label = torch.randint(0, 14, (2, 9, 13), generator=g)             # 13 == C appears: one_hot(label, C+1) accepts it
    assert torch.equal(st.rce(logits, label.clone()), O.reverse_cross_entropy(logits, label.clone(), 13, 12))
    with pytest.raises(RuntimeError):
        st.rce(logits, torch.full((2, 9, 13), 14))
    ok = label.clamp(max=12)
    assert torch.equal(st.seg_loss(logits, ok), O.cross_entropy(logits, ok, 12))
    for th in (None, 0.4):
        vecs, ids = st.calculate_mean_vector(feat, logits, thresh=th)
        ovecs, oids = O.class_mean_vectors(feat, logits, 13, thresh=th, overflow_channel=12, use_conf_mask=True)
        assert ids == oids and all(torch.equal(a, b) for a, b in zip(vecs, ovecs))


def test_eval_and_losses_live(ref):
    r = np.random.RandomState(33)
    for c in (13, 19):
        gt = r.randint(0, c + 1, size=(3, 20, 30)).astype(np.int64)
        pred = r.randint(0, c, size=(3, 20, 30)).astype(np.int64)
        a = ref.utils_eval_tools.evaluate(list(pred), list(gt), c)
        b = O.evaluate(list(pred), list(gt), c)
        assert all(x == y for x, y in zip(a[:4], b[:4]))
        assert np.array_equal(ref.utils_eval_tools._fast_hist(pred.flatten(), gt.flatten(), c), O.fast_hist(pred.flatten(), gt.flatten(), c))
    g = torch.Generator().manual_seed(34)
    x = torch.randn(2, 13, 6, 7, generator=g)
    y = torch.randint(0, 13, (2, 6, 7), generator=g)
    assert torch.equal(ref.utils_loss.SemanticConsistency((2, 4))(x, y.clone()), O.semantic_consistency(x, y.clone(), (2, 4)))
    assert torch.equal(ref.utils_loss.FocalLoss(gamma=2)(x, y), O.focal_loss(x, y, 2))
