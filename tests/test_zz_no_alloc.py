"""tss_class_means_noalloc: the class-mean op (labels -> segmented sums -> means) with its scratch in a static device
array inside the library -- no cudaMalloc, no caller workspace.  It has no counterpart in the reference;
the check is therefore against the existing entry points
tss_proto_labels / tss_proto_accumulate / tss_proto_finalize on the same inputs, on two calls in a row (the array keeps
the first call's contents), called straight through the C ABI.  Collected last on purpose (file name).
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 4 << 20        # TSS_NOALLOC_SCRATCH_BYTES of include/tss_b200.h
RTOL_SUMS = 2e-6               # fp32 RED.ADD order differs between two launches


@pytest.fixture(scope="module")
def api():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import ops
    from b200seg._lib import lib
    return ops, lib()


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _outputs(n, c, k, dev):
    return dict(count_all=torch.full((n, c), -7, dtype=torch.int32, device=dev),
                count_gated=torch.full((n, c), -7, dtype=torch.int32, device=dev),
                vectors=torch.full((n, c, k), float("nan"), dtype=torch.float32, device=dev),
                valid=torch.full((n, c), 77, dtype=torch.uint8, device=dev),
                vector_sum=torch.full((n, c), float("nan"), dtype=torch.float32, device=dev))


def _alloc_counters():
    # cumulative counts (they never go down when earlier tests' tensors are garbage-collected during the call)
    st = torch.cuda.memory_stats()
    return st.get("allocation.all.allocated", 0), st.get("segment.all.allocated", 0)


def _noalloc(L, feat, logits, c, min_pixels=10):
    n, k = feat.shape[0], feat.shape[1]
    plane = feat.shape[2] * feat.shape[3]
    o = _outputs(n, c, k, feat.device)
    torch.cuda.synchronize()
    before = _alloc_counters()
    rc = L.tss_class_means_noalloc(_ptr(feat), _ptr(logits), n, c, k, plane, min_pixels, _ptr(o["count_all"]),
                                   _ptr(o["count_gated"]), _ptr(o["vectors"]), _ptr(o["valid"]), _ptr(o["vector_sum"]),
                                   _stream())
    assert rc == 0, L.tss_last_error_string().decode()
    torch.cuda.synchronize()
    assert _alloc_counters() == before   # no block handed out by torch's allocator, no cudaMalloc of a new segment
    return o


def _existing(L, feat, logits, c, min_pixels=10):
    """the same op through the existing entry points, scratch in caller (torch) memory"""
    n, k = feat.shape[0], feat.shape[1]
    plane = feat.shape[2] * feat.shape[3]
    dev = feat.device
    rec = torch.empty(max(int(L.tss_proto_records_bytes(n, plane)), 16), dtype=torch.uint8, device=dev)
    seg = torch.empty((n, plane), dtype=torch.uint8, device=dev)
    sums = torch.empty((n, c, k), dtype=torch.float32, device=dev)
    o = _outputs(n, c, k, dev)
    st = _stream()
    assert L.tss_proto_labels(_ptr(logits), None, n, c, plane, c, 0, -1.0, _ptr(seg), _ptr(rec), _ptr(sums), sums.numel(), st) == 0
    assert L.tss_proto_accumulate(_ptr(feat), _ptr(rec), n, k, plane, c, _ptr(sums), 0, 0, st) == 0
    assert L.tss_proto_finalize(_ptr(sums), _ptr(rec), n, c, k, plane, min_pixels, _ptr(o["count_all"]), _ptr(o["count_gated"]),
                                _ptr(o["vectors"]), _ptr(o["valid"]), _ptr(o["vector_sum"]), st) == 0
    torch.cuda.synchronize()
    return o


def _same(a, b):
    for name in ("count_all", "count_gated", "valid"):
        assert torch.equal(a[name], b[name]), name
    m = a["valid"].bool()
    va, vb = a["vectors"][m].double().cpu().numpy(), b["vectors"][m].double().cpu().numpy()
    assert np.isfinite(va).all()
    assert np.abs(va - vb).max() <= RTOL_SUMS * max(np.abs(vb).max(), 1e-30)
    sa, sb = a["vector_sum"][m].double().cpu().numpy(), b["vector_sum"][m].double().cpu().numpy()
    # a vector's sum adds K values of that accuracy
    assert np.abs(sa - sb).max() <= RTOL_SUMS * max(np.abs(vb).max(), 1e-30) * a["vectors"].shape[2]


def _inputs(seed, n, c, k, h, w, rare=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    logits = 3.0 * torch.randn(n, c, h, w, generator=g, device="cuda")
    if rare:                       # a class with fewer than min_pixels pixels and one that never wins
        logits[:, 3] = -50.0
        logits[:, 5] = -50.0
        logits[:, 5, 0, :4] = 50.0
    feat = torch.randn(n, k, h, w, generator=g, device="cuda")
    return feat, logits


def test_two_calls_in_a_row_match_the_existing_entry_points(api):
    ops, L = api
    # first call: the reference's CPU-runnable shape (4 x 13 x 41 x 81, K = 256)
    f1, l1 = _inputs(9101, 4, 13, 256, 41, 81)
    _same(_noalloc(L, f1, l1, 13), _existing(L, f1, l1, 13))
    # second call, straight after, different extents and contents: the array still holds the first call's records and
    # sums, so anything stale would show (odd plane, rare / absent classes, 19 classes, K = 64)
    f2, l2 = _inputs(9102, 3, 19, 64, 33, 65, rare=True)
    got = _noalloc(L, f2, l2, 19)
    ref = _existing(L, f2, l2, 19)
    _same(got, ref)
    assert not got["valid"][:, 3].any() and not got["valid"][:, 5].any()
    # ... and the first inputs again give the first answer
    _same(_noalloc(L, f1, l1, 13), _existing(L, f1, l1, 13))
    # the host mirror of the existing path agrees too (ops.class_means is what the drop-in calls)
    m = ops.class_means(f1, l1, 13)
    again = _noalloc(L, f1, l1, 13)
    assert torch.equal(again["valid"], m.valid) and torch.equal(again["count_all"], m.count_all)


def test_largest_batch_that_fits_and_rejection_beyond_it(api):
    _, L = api
    # cfg2's batch (16 x 13 x 256 x 81 x 241) needs 1.63 MB: fits
    f, l = _inputs(9103, 16, 13, 256, 81, 241)
    _same(_noalloc(L, f, l, 13), _existing(L, f, l, 13))
    # 48 such images need 4.9 MB of scratch: rejected before anything is launched, outputs untouched
    n, c, k, h, w = 48, 13, 256, 81, 241
    feat = torch.zeros(n, k, h, w, device="cuda")
    logits = torch.zeros(n, c, h, w, device="cuda")
    o = _outputs(n, c, k, feat.device)
    rc = L.tss_class_means_noalloc(_ptr(feat), _ptr(logits), n, c, k, h * w, 10, _ptr(o["count_all"]), _ptr(o["count_gated"]),
                                   _ptr(o["vectors"]), _ptr(o["valid"]), _ptr(o["vector_sum"]), _stream())
    assert rc != 0
    msg = L.tss_last_error_string().decode()
    assert "scratch" in msg and str(SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert int(o["valid"][0, 0]) == 77 and int(o["count_all"][0, 0]) == -7
    # null pointers and a channel count the accumulate kernel cannot take are refused as well
    assert L.tss_class_means_noalloc(None, _ptr(logits), 1, c, k, h * w, 10, _ptr(o["count_all"]), _ptr(o["count_gated"]),
                                     _ptr(o["vectors"]), _ptr(o["valid"]), _ptr(o["vector_sum"]), _stream()) != 0
    assert L.tss_class_means_noalloc(_ptr(feat), _ptr(logits), 1, c, 48, h * w, 10, _ptr(o["count_all"]), _ptr(o["count_gated"]),
                                     _ptr(o["vectors"]), _ptr(o["valid"]), _ptr(o["vector_sum"]), _stream()) != 0
