"""Pseudo-label files: what b200seg.formats writes is what the reference's own reader code
loads (data/freiburg_dataset.py:108-125), and what the reference's writer code produces
(generate_pseudo_label.py:78,91,94) is what b200seg.formats reads."""
import os

import numpy as np
import pytest
import torch
from PIL import Image

from oracle import label_path as O


def _case(seed, n=3, h=40, w=96, c=13):
    r = np.random.RandomState(seed)
    label = r.randint(0, c, (n, h, w)).astype(np.uint8)
    conf = r.uniform(0.05, 1.0, (n, h, w)).astype(np.float16)
    names = ["/data/seq_%02d/fl_ir_aligned_%04d.png" % (seed, k) for k in range(n)]
    return label, conf, names


def test_writers_produce_reference_readable_files(tmp_path):
    from b200seg import formats
    label, conf, names = _case(1)
    d = str(tmp_path / "hard")
    formats.save_hard_batch(d, names, torch.from_numpy(label), torch.from_numpy(conf))
    for k, name in enumerate(names):
        base = os.path.basename(name)
        # the reference's reader lines, verbatim semantics (freiburg_dataset.py:113-118)
        got = np.array(Image.open(os.path.join(d, base)), dtype=np.uint8)
        got_conf = np.load(os.path.join(d, base.replace(".png", "_conf.npy")))
        assert got.dtype == np.uint8 and (got == label[k]).all()
        assert got_conf.dtype == np.float16 and (got_conf == conf[k]).all()
    prob = torch.softmax(torch.randn(3, 13, 11, 21), 1)
    d2 = str(tmp_path / "soft")
    formats.save_soft_batch(d2, names, prob)
    for k, name in enumerate(names):
        got = formats.read_soft(d2, name)
        assert got.dtype == np.float32 and got.shape == (13, 11, 21) and (got == prob[k].numpy()).all()


def test_read_hard_files_reads_reference_written_files(tmp_path):
    from b200seg import formats
    label, conf, names = _case(2)
    d = str(tmp_path)
    for k, name in enumerate(names):  # the reference's writer lines (generate_pseudo_label.py:91,94)
        base = os.path.basename(name)
        Image.fromarray(label[k]).save(os.path.join(d, base))
        np.save(os.path.join(d, base.replace(".png", "_conf.npy")), conf[k].astype(np.float32).astype(np.float16))
    for k, name in enumerate(names):
        lab, cf = formats.read_hard_files(d, name)
        assert (lab == label[k]).all() and cf.dtype == np.float16 and (cf == conf[k]).all()


@pytest.mark.gpu
def test_load_hard_batch_matches_reference_reader(tmp_path):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import formats
    label, conf, names = _case(3, n=4, h=64, w=160)
    conf[0, :4, :4] = np.float16(0.3)           # exactly on the threshold: fp16(0.3) <= fp16(0.3) rejects
    d = str(tmp_path)
    formats.save_hard_batch(d, names, torch.from_numpy(label).cuda(), torch.from_numpy(conf).cuda())
    size_wh = (700, 320)                         # image.size of the Freiburg crop (W, H)
    got = formats.load_hard_batch(d, names, 13, "cuda", threshold=0.3, ignore_index=12, keep=(5, 11), size_wh=size_wh)
    for k, name in enumerate(names):
        lab, cf = formats.read_hard_files(d, name)
        ref = O.threshold_hard_label(lab, cf, 0.3, 12, keep=(5, 11))
        ref = np.array(Image.fromarray(ref).resize(size_wh, Image.NEAREST), dtype=np.uint8)   # freiburg_dataset.py:124
        assert (got[k].cpu().numpy() == ref).all()
    raw = formats.load_hard_batch(d, names, 13, "cuda", threshold=None)
    assert (raw.cpu().numpy() == label).all()


@pytest.mark.gpu
def test_load_soft_batch_matches_loader_resize(tmp_path):
    """soft .npy maps -> device -> the loader's bilinear resize (utils/self_train_augmentation.py:21-24)"""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import formats
    from oracle import label_path as O
    r = np.random.RandomState(5)
    names = ["a_%d.png" % i for i in range(3)]
    maps = [r.dirichlet(np.ones(13), size=(9, 17)).transpose(2, 0, 1).astype(np.float32) for _ in names]
    for n, m in zip(names, maps):
        np.save(os.path.join(str(tmp_path), n.replace(".png", ".npy")), m)
    out = formats.load_soft_batch(str(tmp_path), names, (320, 700), "cuda")
    ref = O.upsample_bilinear_ac(torch.from_numpy(np.stack(maps)), (320, 700))
    assert tuple(out.shape) == (3, 13, 320, 700)
    assert (out.cpu() - ref).abs().max().item() <= 2e-6
