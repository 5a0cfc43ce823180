"""CPU-side checks: the C-ABI library loads and exports every symbol include/tss_b200.h declares (no compute
calls), the host-side helpers agree with the oracle, and the product path refuses CPU tensors."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import label_path as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "tss_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tss_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_tables_agree():
    from b200seg import _lib
    assert sorted(_lib.EXPORTED) == _header_symbols()


def test_library_exports_every_declared_symbol():
    from b200seg import _build
    path = _build.build()
    handle = ctypes.CDLL(path)
    for name in _header_symbols():
        assert hasattr(handle, name), name
    handle.tss_version.restype = ctypes.c_int
    assert handle.tss_version() == 100
    handle.tss_proto_records_bytes.restype = ctypes.c_size_t
    handle.tss_proto_records_bytes.argtypes = [ctypes.c_int64, ctypes.c_int64]
    assert handle.tss_proto_records_bytes(2, 1000) == 2 * 2 * 1776   # TSS_PROTO_RECORD_BYTES
    assert handle.tss_proto_records_bytes(0, 1000) == 0


def test_argument_rejection_needs_no_gpu():
    from b200seg import _lib
    L = _lib.lib()
    assert L.tss_confusion(None, None, 0, 10, 13, None, None, 0, None) == 1          # TSS_E_ARG: null hist
    assert b"null" in L.tss_last_error_string()
    buf = ctypes.create_string_buffer(64)
    p = ctypes.cast(buf, ctypes.c_void_p)
    assert L.tss_confusion(p, p, 0, 10, 99, p, None, 0, None) == 2                   # TSS_E_CLASSES
    assert L.tss_confusion(p, p, 7, 10, 13, p, None, 0, None) == 1                   # bad dtype
    assert L.tss_proto_accumulate(p, p, 1, 33, 10, 13, p, 1, 0, None) == 1              # channels % 32
    assert L.tss_confusion(p, p, 0, 0, 13, p, None, 0, None) == 0                    # empty input is a no-op
    # fused pseudo-label + confusion matrix: gt and hist are required, flags are validated
    assert L.tss_pseudo_label_hist(p, None, 1, 13, 4, 4, 8, 8, 0, 0, p, None, None, None, None, None, 12, None, p, 0, None) == 1
    assert b"gt" in L.tss_last_error_string()
    assert L.tss_pseudo_label_hist(p, None, 1, 13, 4, 4, 8, 8, 0, 0, p, None, None, None, None, None, 12, p, p, 0x80, None) == 1
    assert L.tss_pseudo_label_hist(p, None, 0, 13, 4, 4, 8, 8, 0, 0, p, None, None, None, None, None, 12, p, p, 0, None) == 0   # no images
    # the no-allocation class-mean entry: rejected before any CUDA call -- null pointer, channels % 32, a batch whose
    # scratch exceeds the built-in 4 MiB array; an empty batch is a no-op
    args = (p, p, p, p, p)
    assert L.tss_class_means_noalloc(None, p, 1, 13, 256, 100, 10, *args, None) == 1
    assert L.tss_class_means_noalloc(p, p, 1, 13, 48, 100, 10, *args, None) == 1
    assert L.tss_class_means_noalloc(p, p, 1, 99, 256, 100, 10, *args, None) == 2
    assert L.tss_class_means_noalloc(p, p, 48, 13, 256, 81 * 241, 10, *args, None) == 1
    assert b"scratch" in L.tss_last_error_string() and str(4 << 20).encode() in L.tss_last_error_string()
    assert L.tss_class_means_noalloc(p, p, 0, 13, 256, 100, 10, *args, None) == 0


def test_cpu_tensors_are_refused():
    from b200seg import ops
    with pytest.raises(ops.TssError):
        ops.confusion_matrix(torch.zeros(8, dtype=torch.uint8), torch.zeros(8, dtype=torch.uint8), 13)
    with pytest.raises(ops.TssError):
        ops.pseudo_label(torch.zeros(1, 13, 4, 4), size=(8, 8))
    with pytest.raises(ops.TssError):
        ops.prototype_distance(torch.zeros(1, 32, 4, 4), torch.zeros(13, 32))


def test_no_product_module_imports_the_oracle():
    pkg = os.path.join(ROOT, "thermal-semantic-segmentation_b200")
    for name in os.listdir(pkg):
        if name.endswith(".py"):
            src = open(os.path.join(pkg, name)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), name


@pytest.mark.parametrize("n_in,n_out", [(41, 320), (81, 640), (33, 256), (88, 700), (700, 88), (5, 5), (1, 7)])
def test_nearest_tables_match_oracle(n_in, n_out):
    from b200seg import ops
    assert (ops.pil_nearest_table(n_in, n_out) == O.pil_nearest_index_np(n_in, n_out)).all()
    lab = torch.arange(n_in, dtype=torch.float32).reshape(1, 1, 1, n_in)
    ref = torch.nn.functional.interpolate(lab, size=(1, n_out), mode="nearest").reshape(-1).long().numpy()
    assert (ops.torch_nearest_table(n_in, n_out) == ref).all()


def test_weak_param_table_matches_reference_arithmetic():
    from b200seg import ops
    wp = {"RandomSized": [torch.tensor([563, 704, 512]), torch.tensor([1126, 1408, 1024])],
          "RandomCrop": [torch.tensor([13, 101, 0]), torch.tensor([269, 357, 256]), torch.tensor([57, 300, 0]),
                         torch.tensor([569, 812, 512])],
          "RandomHorizontallyFlip": [torch.tensor(True), torch.tensor(False), torch.tensor(True)]}
    table, extents = ops.weak_param_table(wp, 3)
    feat = torch.randn(3, 4, 40, 80)
    ref = O.full2weak(feat, wp)
    assert extents == [tuple(ref.shape[2:])] * 3
    for i in range(3):
        hh, ww = int(wp["RandomSized"][0][i]), int(wp["RandomSized"][1][i])
        assert table[i, 0] == int(hh / 4) and table[i, 1] == int(ww / 4)
        assert table[i, 2] == int(int(wp["RandomCrop"][0][i]) / 4) and table[i, 3] == int(int(wp["RandomCrop"][2][i]) / 4)
        assert table[i, 4] == int(bool(wp["RandomHorizontallyFlip"][i]))


@pytest.mark.parametrize("c", [13, 19])
def test_metrics_from_hist_is_the_reference_arithmetic(c):
    from b200seg import dropin
    r = np.random.RandomState(5)
    hist = r.randint(0, 1000, size=(c, c)).astype(np.int64)
    hist[3] = 0                      # a class absent from the ground truth -> nan in acc_cls / iu
    a, b = dropin.metrics_from_hist(hist, c), O.metrics_from_hist(hist, c)
    for x, y in zip(a[:4], b[:4]):
        assert (np.isnan(x) and np.isnan(y)) or x == y
    assert list(a[4].keys()) == list(b[4].keys())
    with pytest.raises(ValueError):
        dropin.metrics_from_hist(np.ones((5, 5)), 5)
    with pytest.raises(ValueError):
        dropin.evaluate([], [], 5)


def test_shard_range_partitions_the_image_order():
    from b200seg import sharding
    for total in (0, 1, 7, 16, 10000, 50000):
        for ws in (1, 2, 3, 8):
            spans = [sharding.shard_range(total, r, ws) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)
