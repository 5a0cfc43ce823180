"""The reference's OWN modules with dropin.install() applied: the scripts' entry points --
generate_pseudo_label.generate_pl, segmentation_evaluate.seg_validate, cal_prototype.Class_Features,
self_training.SelfTrain.step and the Freiburg dataset's threshold block -- are called through the reference's module
namespaces, unmodified, and land in the CUDA kernels.  Results are checked against the golden fixtures (produced by
the same functions before rebinding).  Needs the reference tree: /root/reference or baseline/_ref/ (staged by
__graft_entry__.build(), travels to the GPU box with the snapshot); skipped when neither is there."""
import os
import types

import numpy as np
import pytest
import torch
from PIL import Image

import inputs as gi
import ref_import
from oracle import label_path as O

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_import.reference_available(), reason="reference tree not present")]


@pytest.fixture(scope="module")
def ref():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import dropin
    ns = ref_import.import_reference()
    originals = types.SimpleNamespace(seg_validate=ns.segmentation_evaluate.seg_validate)
    dropin.install({"cal_prototype": ns.cal_prototype, "self_training": ns.self_training,
                    "generate_pseudo_label": ns.generate_pseudo_label, "segmentation_evaluate": ns.segmentation_evaluate,
                    "data.freiburg_dataset": ns.data_freiburg_dataset, "data": ns.data,
                    "utils": ns.utils, "utils.eval_tools": ns.utils_eval_tools, "utils.loss": ns.utils_loss})
    ns.originals = originals
    ns.dropin = dropin
    return ns


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


class _ReplayNet(torch.nn.Module):
    def __init__(self, seq):
        super().__init__()
        self.seq, self.i = seq, 0

    def forward(self, images):
        out = self.seq[self.i % len(self.seq)]
        self.i += 1
        return {"out": out.to(images.device).clone()}


def test_install_rebinds_every_entry_point(ref):
    d = ref.dropin
    assert ref.cal_prototype.Class_Features is d.Class_Features
    assert ref.self_training.SelfTrain is d.SelfTrain
    assert ref.generate_pseudo_label.generate_pl is d.generate_pl
    assert ref.segmentation_evaluate.seg_validate is d.seg_validate
    assert ref.segmentation_evaluate.evaluate is d.evaluate and ref.utils_eval_tools._fast_hist is d._fast_hist
    assert ref.utils_loss.SemanticConsistency is d.SemanticConsistency and ref.utils_loss.FocalLoss is d.FocalLoss
    assert getattr(ref.data_freiburg_dataset.Freiburg.__init__, "_b200seg_hooked", False)


@pytest.mark.parametrize("branch", ["soft", "hard", "flip"])
def test_reference_generate_pl_after_install(ref, golden, tmp_path, branch):
    name = "pl_small"
    n, c, h, w, H, W, _ = gi.PL_CASES[name]
    logits, flip_logits = gi.pl_logits(name)
    names = ["img%02d.png" % i for i in range(n)]
    batch = [{"image": torch.zeros(n, 1, H, W), "img_path": names}]
    args = types.SimpleNamespace(root=str(tmp_path), pseudo_type=branch, checkpoint_name="ck.pth", soft=(branch == "soft"),
                                 flip=(branch == "flip"))
    with torch.no_grad():
        ref.generate_pseudo_label.generate_pl(_ReplayNet([T(logits), T(flip_logits)]), batch, torch.device("cuda"), args)
    d = os.path.join(str(tmp_path), "pseudo_labels", branch, "ck")
    if branch == "soft":
        arr = np.stack([np.load(os.path.join(d, nm.replace(".png", ".npy"))) for nm in names])
        np.testing.assert_allclose(arr, golden[name + "_soft_prob"], rtol=2e-6, atol=1e-9)
        return
    lab = np.stack([np.array(Image.open(os.path.join(d, nm))) for nm in names])
    conf = np.stack([np.load(os.path.join(d, nm.replace(".png", "_conf.npy"))) for nm in names])
    assert lab.dtype == np.uint8 and conf.dtype == np.float16
    ref_lab, ref_conf = golden["%s_%s_label" % (name, branch)], golden["%s_%s_conf" % (name, branch)]
    if branch == "hard":
        prob = golden[name + "_soft_prob"]
    else:
        a = O.upsample_bilinear_ac(O.pseudo_label_soft(T(logits)), (H, W))
        b = O.upsample_bilinear_ac(O.pseudo_label_soft(T(flip_logits)), (H, W))
        prob = torch.softmax((a + torch.flip(b, dims=[1])) / 2, 1).numpy()
    srt = np.sort(prob, axis=1)
    safe = (srt[:, -1] - srt[:, -2]) >= 1e-6
    assert (lab[safe] == ref_lab[safe]).all()
    ulp = np.abs(conf.view(np.int16).astype(np.int32) - ref_conf.view(np.int16).astype(np.int32))
    assert ulp.max() <= 1 and (ulp > 0).mean() < 0.005


def test_reference_seg_validate_after_install(ref):
    """seg_validate through the rebinding vs the reference's original function on the CPU, same stand-in network and
    loader: identical mIoU / per-class IoU (bit-identical float64 given identical labels), loss within 1e-5."""
    r = gi.rs(3100)
    nb, bs, c = 3, 4, 13
    logits = [T((3 * r.standard_normal((bs, c, 33, 65))).astype(np.float32)) for _ in range(nb)]
    labels = [T(r.randint(0, 13, size=(bs, 256, 512)).astype(np.int64)) for _ in range(nb)]
    loader = [{"image": torch.zeros(bs, 1, 256, 512), "label": lab} for lab in labels]
    args = types.SimpleNamespace(visualize_prediction=None)
    loss_cpu = torch.nn.CrossEntropyLoss(ignore_index=12, reduction="mean")
    m_ref, l_ref, iu_ref = ref.originals.seg_validate(args, _ReplayNet(logits), loader, loss_cpu, torch.device("cpu"), None, 13)
    net = _ReplayNet(logits)
    m_got, l_got, iu_got = ref.segmentation_evaluate.seg_validate(args, net, loader, loss_cpu, torch.device("cuda"), None, 13)
    # predictions can differ only at upsampled-logit near-ties (< 1e-6); with 3 * randn logits none is expected
    ups = [O.upsample_bilinear_ac(x, (256, 512)) for x in logits]
    tight = min(float((u.topk(2, dim=1)[0][:, 0] - u.topk(2, dim=1)[0][:, 1]).min()) for u in ups)
    if tight > 1e-6:
        assert m_got == m_ref and all(iu_got[k] == iu_ref[k] or np.isnan(iu_ref[k]) for k in iu_ref)
    else:
        assert abs(m_got - m_ref) < 1e-5
    np.testing.assert_allclose(l_got, l_ref, rtol=1e-5)


def test_reference_class_features_loop_after_install(ref, golden):
    """The loop body of cal_prototype.calc_prototype (:71-75) written against the reference's module namespace."""
    cp = ref.cal_prototype
    feat, logits, _ = gi.proto_inputs()
    cf = cp.Class_Features(torch.device("cuda"), numbers=13)
    for f in (feat, gi.randn(302, *feat.shape)):
        out = {"feat": T(f).cuda(), "out": T(logits).cuda()}
        vectors, ids = cf.calculate_mean_vector(out["feat"], out["out"])
        for t in range(len(ids)):
            cf.update_objective_SingleVector(ids[t], vectors[t].detach().cpu(), "mean")
    assert np.array_equal(cf.objective_vectors_num.cpu().numpy(), golden["cf_objective_num"])
    got, want = cf.objective_vectors.cpu().numpy(), golden["cf_objective_vectors"]
    assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max()


def test_reference_freiburg_threshold_block_after_install(ref, golden, tmp_path):
    """Freiburg(..., self_train=True) built through the reference's class after install(): the stored pseudo-labels
    are thresholded / whitelisted / resized on the device when the dataset is constructed, and the unmodified
    __getitem__ hands the augmentations exactly the label map the reference's numpy block produces."""
    fd = ref.data_freiburg_dataset
    label, conf = gi.threshold_inputs()
    tmp = str(tmp_path)
    for sub in ("image_list", "lp", "soft"):
        os.makedirs(os.path.join(tmp, sub))
    img_path = os.path.join(tmp, "fr_0001.png")
    Image.fromarray((np.full((320, 960), 23000)).astype(np.uint16)).save(img_path)
    lab_path = os.path.join(tmp, "fr_0001_label.png")
    Image.fromarray(np.zeros((320, 960), np.uint8)).save(lab_path)
    with open(os.path.join(tmp, "image_list", "train_IR_data.txt"), "w") as f:
        f.write(img_path + "\n")
    with open(os.path.join(tmp, "image_list", "train_IR_label.txt"), "w") as f:
        f.write(lab_path + "\n")
    Image.fromarray(label).save(os.path.join(tmp, "lp", "fr_0001.png"))
    np.save(os.path.join(tmp, "lp", "fr_0001_conf.npy"), conf.astype(np.float16))
    np.save(os.path.join(tmp, "soft", "fr_0001.npy"), np.zeros((13, 4, 4), np.float32))
    captured = {}

    def capture_aug(image, lab, label_hard, label_soft):
        captured["hard"] = np.array(label_hard).copy()
        raise StopIteration

    args = types.SimpleNamespace(proto_rectify=True, path_soft=os.path.join(tmp, "soft"), path_lp=os.path.join(tmp, "lp"),
                                 threshold=0.3, ignore_index=12, num_classes=13)
    ds = fd.Freiburg(args, root=tmp, split="train", domain="IR", with_label=True, augmentations=capture_aug, self_train=True)
    assert ds.args.threshold is None and ds.args.path_lp.endswith("_b200thr") and args.threshold == 0.3
    try:
        ds[0]
    except StopIteration:
        pass
    assert np.array_equal(captured["hard"], golden["thr_label_hard_700x320"])
