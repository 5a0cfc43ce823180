"""Tensor-level host API over the C ABI (include/tss_b200.h).

PyTorch is plumbing here: it owns device memory and the stream; every piece of label arithmetic runs in
the hand-written sm_100a kernels of libtss_b200.so.  CPU tensors are rejected -- there is no fallback.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import TssError, check, lib  # noqa: F401

MAX_CLASSES = 32
PL_UP_SOFTMAX, PL_UP_ARGMAX, PL_PROB_UP, PL_PROB_UP_SOFTMAX = 0, 1, 2, 3
FLIP_NONE, FLIP_CLASS, FLIP_WIDTH = 0, 1, 2
UPDATE_MEAN, UPDATE_MOVING_AVERAGE = 0, 1
HIST_AUTO, HIST_ATOMIC, HIST_PACKED = 0, 1, 2
PL_FLAG_FORCE_TILE = 1
DIST_DIRECT, DIST_EXPANDED = 1, 2


PL_FLAG_ROWS2 = 2
PL_FLAG_REGS96 = 4


def pl_flags(force_tile: bool = False, sm_warps: int = 0, cta_warps: int = 0, waves: int = 0, rows2: bool = False,
             regs96: bool = False) -> int:
    """`flags` of tss_pseudo_label: A/B switch of the tile kernel; warps per SM, warps per CTA and number of waves of
    the register-resident kernel (co-resident forms)."""
    return (PL_FLAG_FORCE_TILE if force_tile else 0) | (PL_FLAG_ROWS2 if rows2 else 0) | (PL_FLAG_REGS96 if regs96 else 0) | \
        ((int(sm_warps) & 0xff) << 8) | \
        ((int(cta_warps) & 0xf) << 16) | ((int(waves) & 0xf) << 20)


# The co-resident step form measured best on B200 at the cfg2 shape (profiles/r02_overlap_sweep.txt): pseudo-label kernel
# in 4-warp CTAs, two waves, up to 16 warps per SM; accumulate kernel with 6 consumer warps on a high-priority stream
CORESIDENT_PL_FLAGS = pl_flags(sm_warps=16, cta_warps=4, waves=2)
CORESIDENT_CONSUMERS = 6


def dist_form(variant: int = 0, prefetch: int = -1, split: int = 0, algo: int = 0) -> int:
    """`form` of tss_proto_distance (TSS_DIST_FORM): 0 everywhere = automatic."""
    return (variant & 0xf) | (((prefetch + 1) & 0xf) << 4) | ((split & 0xf) << 8) | ((algo & 0xf) << 12)
_DT = {torch.uint8: 0, torch.int32: 1, torch.int64: 2}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dev(t: torch.Tensor, dtype=None, name: str = "tensor") -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s: expected a torch.Tensor, got %s" % (name, type(t).__name__))
    if not t.is_cuda:
        raise TssError("%s is on %s: b200seg runs on CUDA only (no CPU fallback)" % (name, t.device))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _label_tensor(t: torch.Tensor, name: str) -> torch.Tensor:
    t = _dev(t, name=name)
    if t.dtype not in _DT:
        t = t.to(torch.int64)
    return t


# ------------------------------------------------------------------------------------------- A14
def confusion_matrix(pred: torch.Tensor, gt: torch.Tensor, num_classes: int,
                     hist: Optional[torch.Tensor] = None, bad: Optional[torch.Tensor] = None, variant: int = HIST_AUTO):
    """hist[C,C] int64 (+= when given), rows = ground truth.  Returns (hist, bad_keys) device tensors."""
    pred, gt = _label_tensor(pred, "pred"), _label_tensor(gt, "gt")
    if pred.dtype != gt.dtype:
        pred, gt = pred.to(torch.int64), gt.to(torch.int64)
    if pred.numel() != gt.numel():
        raise ValueError("pred and gt differ in size: %d vs %d" % (pred.numel(), gt.numel()))
    if hist is None:
        hist = torch.zeros(num_classes, num_classes, dtype=torch.int64, device=pred.device)
    if bad is None:
        bad = torch.zeros(1, dtype=torch.int64, device=pred.device)
    check(lib().tss_confusion(_ptr(pred), _ptr(gt), _DT[pred.dtype], pred.numel(), num_classes, _ptr(hist),
                              _ptr(bad), int(variant), _stream()), "tss_confusion")
    return hist, bad


# ------------------------------------------------------------------------------------------- A1-A4
def softmax_nchw(logits: torch.Tensor) -> torch.Tensor:
    x = _dev(logits, torch.float32, "logits")
    n, c = x.shape[0], x.shape[1]
    plane = x[0, 0].numel()
    out = torch.empty_like(x)
    check(lib().tss_softmax_nchw(_ptr(x), _ptr(out), n, c, plane, _stream()), "tss_softmax_nchw")
    return out


def _host_tables(num_classes, thr, keep):
    thr_arr = keep_arr = None
    if thr is not None:
        thr_np = np.asarray(thr, dtype=np.float32)
        if thr_np.ndim == 0:
            thr_np = np.full(num_classes, float(thr_np), np.float32)
        if thr_np.size != num_classes:
            raise ValueError("thr must be a scalar or have num_classes entries")
        thr_arr = (ctypes.c_float * num_classes)(*thr_np.tolist())
    if keep is not None:
        mask = np.zeros(num_classes, np.uint8)
        k = np.asarray(list(keep))
        if k.dtype == np.bool_ and k.size == num_classes:
            mask[:] = k
        else:
            mask[k.astype(np.int64)] = 1
        keep_arr = (ctypes.c_uint8 * num_classes)(*mask.tolist())
    return thr_arr, keep_arr


@dataclass
class PseudoLabels:
    label: Optional[torch.Tensor]       # [N,H,W] uint8 argmax
    conf16: Optional[torch.Tensor]      # [N,H,W] float16 confidence as the reference stores it
    conf32: Optional[torch.Tensor]      # [N,H,W] float32 confidence before the storage cast
    label_thr: Optional[torch.Tensor]   # [N,H,W] uint8 after the class-wise threshold / whitelist


def pseudo_label(src: torch.Tensor, size: Optional[Sequence[int]] = None, mode: int = PL_UP_SOFTMAX,
                 src_flip: Optional[torch.Tensor] = None, flip: int = FLIP_NONE, thr=None, keep=None,
                 ignore_index: int = 255, want_label: bool = True, want_conf16: bool = True,
                 want_conf32: bool = False, want_thr: Optional[bool] = None, flags: int = 0,
                 gt: Optional[torch.Tensor] = None, hist: Optional[torch.Tensor] = None) -> PseudoLabels:
    """Fused upsample -> softmax -> argmax/confidence -> storage casts -> threshold (tss_pseudo_label).
    With ``gt`` ([N,H,W] uint8) and ``hist`` ([C,C] int64, accumulated in place) the confusion matrix of (gt, label) is
    taken in the same pass (tss_pseudo_label_hist)."""
    x = _dev(src, torch.float32, "src")
    if x.dim() != 4:
        raise ValueError("src must be [N,C,h,w]")
    n, c, h, w = x.shape
    H, W = (h, w) if size is None else (int(size[0]), int(size[1]))
    x2 = None
    if flip != FLIP_NONE:
        if src_flip is None:
            raise ValueError("flip requested without src_flip")
        x2 = _dev(src_flip, torch.float32, "src_flip")
        if x2.shape != x.shape:
            raise ValueError("src_flip shape differs from src")
    if want_thr is None:
        want_thr = thr is not None or keep is not None
    dev = x.device
    label = torch.empty((n, H, W), dtype=torch.uint8, device=dev) if want_label else None
    conf16 = torch.empty((n, H, W), dtype=torch.float16, device=dev) if want_conf16 else None
    conf32 = torch.empty((n, H, W), dtype=torch.float32, device=dev) if want_conf32 else None
    label_thr = torch.empty((n, H, W), dtype=torch.uint8, device=dev) if want_thr else None
    thr_arr, keep_arr = _host_tables(c, thr, keep)
    if (gt is None) != (hist is None):
        raise ValueError("gt and hist go together")
    if gt is not None:
        if gt.dtype != torch.uint8 or hist.dtype != torch.int64 or not hist.is_contiguous():
            raise ValueError("gt must be uint8 and hist a contiguous int64 tensor (it is accumulated in place)")
        g = _dev(gt, torch.uint8, "gt")
        hh = _dev(hist, torch.int64, "hist")
        if g.numel() != n * H * W or hh.numel() != c * c:
            raise ValueError("gt must be [N,H,W] and hist [C,C]")
        check(lib().tss_pseudo_label_hist(_ptr(x), _ptr(x2), n, c, h, w, H, W, mode, flip, _ptr(label), _ptr(conf16),
                                          _ptr(conf32), _ptr(label_thr),
                                          ctypes.cast(thr_arr, ctypes.c_void_p) if thr_arr is not None else None,
                                          ctypes.cast(keep_arr, ctypes.c_void_p) if keep_arr is not None else None,
                                          int(ignore_index), _ptr(g), _ptr(hh), int(flags), _stream()), "tss_pseudo_label_hist")
        return PseudoLabels(label, conf16, conf32, label_thr)
    check(lib().tss_pseudo_label(_ptr(x), _ptr(x2), n, c, h, w, H, W, mode, flip, _ptr(label), _ptr(conf16),
                                 _ptr(conf32), _ptr(label_thr),
                                 ctypes.cast(thr_arr, ctypes.c_void_p) if thr_arr is not None else None,
                                 ctypes.cast(keep_arr, ctypes.c_void_p) if keep_arr is not None else None,
                                 int(ignore_index), int(flags), _stream()), "tss_pseudo_label")
    return PseudoLabels(label, conf16, conf32, label_thr)


def upsample_bilinear(x: torch.Tensor, size: Sequence[int]) -> torch.Tensor:
    """Bilinear, align_corners=True, of the last two dims of a contiguous fp32 tensor."""
    x = _dev(x, torch.float32, "x")
    h, w = x.shape[-2], x.shape[-1]
    H, W = int(size[0]), int(size[1])
    out = torch.empty(x.shape[:-2] + (H, W), dtype=torch.float32, device=x.device)
    planes = x.numel() // (h * w) if h * w else 0
    check(lib().tss_upsample_bilinear(_ptr(x), _ptr(out), planes, h, w, H, W, _stream()), "tss_upsample_bilinear")
    return out


def upsample_bilinear_adjoint(grad_out: torch.Tensor, in_hw: Sequence[int]) -> torch.Tensor:
    """Backward of upsample_bilinear: grad wrt the low-resolution input."""
    g = _dev(grad_out, torch.float32, "grad_out")
    H, W = g.shape[-2], g.shape[-1]
    h, w = int(in_hw[0]), int(in_hw[1])
    out = torch.empty(g.shape[:-2] + (h, w), dtype=torch.float32, device=g.device)
    planes = g.numel() // (H * W) if H * W else 0
    check(lib().tss_upsample_bilinear_adjoint(_ptr(g), _ptr(out), planes, h, w, H, W, _stream()),
          "tss_upsample_bilinear_adjoint")
    return out


def threshold_labels(label: torch.Tensor, conf16: Optional[torch.Tensor], num_classes: int, thr=None, keep=None,
                     ignore_index: int = 255) -> torch.Tensor:
    lab = _dev(label, torch.uint8, "label")
    cf = None if conf16 is None else _dev(conf16, torch.float16, "conf16")
    out = torch.empty_like(lab)
    thr_arr, keep_arr = _host_tables(num_classes, thr, keep)
    check(lib().tss_threshold_labels(_ptr(lab), _ptr(cf), lab.numel(), num_classes,
                                     ctypes.cast(thr_arr, ctypes.c_void_p) if thr_arr is not None else None,
                                     ctypes.cast(keep_arr, ctypes.c_void_p) if keep_arr is not None else None,
                                     int(ignore_index), _ptr(out), _stream()), "tss_threshold_labels")
    return out


def pil_nearest_table(n_in: int, n_out: int) -> np.ndarray:
    """Pillow NEAREST index rule: start 0.5*scale, accumulate += scale in double, truncate
    (libImaging/Geometry.c ImagingScaleAffine; data/freiburg_dataset.py:124)."""
    scale = float(n_in) / float(n_out)
    pos = scale * 0.5
    out = np.empty(n_out, np.int32)
    for i in range(n_out):
        out[i] = int(pos)
        pos += scale
    return np.minimum(out, n_in - 1).astype(np.int32)


def torch_nearest_table(n_in: int, n_out: int) -> np.ndarray:
    """ATen 'nearest' rule: floor(dst * float(in)/out) in fp32 (self_training.py:103)."""
    scale = np.float32(n_in) / np.float32(n_out)
    idx = np.floor(np.arange(n_out, dtype=np.float32) * scale).astype(np.int64)
    return np.minimum(idx, n_in - 1).astype(np.int32)


def resize_nearest_u8(label: torch.Tensor, size_hw: Sequence[int], rule: str = "pil") -> torch.Tensor:
    lab = _dev(label, torch.uint8, "label")
    h, w = lab.shape[-2], lab.shape[-1]
    H, W = int(size_hw[0]), int(size_hw[1])
    mk = pil_nearest_table if rule == "pil" else torch_nearest_table
    ytab = torch.from_numpy(mk(h, H)).to(lab.device)
    xtab = torch.from_numpy(mk(w, W)).to(lab.device)
    out = torch.empty(lab.shape[:-2] + (H, W), dtype=torch.uint8, device=lab.device)
    planes = lab.numel() // (h * w)
    check(lib().tss_gather_nearest_u8(_ptr(lab), _ptr(out), planes, h, w, H, W, _ptr(ytab), _ptr(xtab), _stream()),
          "tss_gather_nearest_u8")
    return out


# ------------------------------------------------------------------------------------------- A5/A6
@dataclass
class ClassMeans:
    """Per-(image, class) masked means of one batch, all on device."""
    vectors: torch.Tensor      # [N,C,K] fp32 (zero where not valid)
    valid: torch.Tensor        # [N,C] uint8: passes the count>0 and count>=10 rules
    vector_sum: torch.Tensor   # [N,C] fp32: sum over K (the reference skips updates whose sum is exactly 0)
    count_all: torch.Tensor    # [N,C] int32 ungated one-hot pixel count
    count_gated: torch.Tensor  # [N,C] int32 gated count (the divisor)
    seg_label: torch.Tensor    # [N,P] uint8 class each pixel was summed into (255 = none)
    sums: torch.Tensor         # [N,C,K] fp32 raw segmented sums


def _check_bank(protos: torch.Tensor, num: torch.Tensor, c: int, k: int) -> None:
    for t, nm in ((protos, "protos"), (num, "num")):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise TssError("%s must be a contiguous fp32 CUDA tensor (updated in place)" % nm)
    if protos.shape != (c, k) or num.numel() != c:
        raise ValueError("protos %s / num %s do not match [%d,%d]" % (tuple(protos.shape), tuple(num.shape), c, k))


@dataclass
class ClassSums:
    """Segmented sums of one batch before the mean / validity step (tss_proto_labels + tss_proto_accumulate)."""
    sums: torch.Tensor         # [N,C,K] fp32
    records: torch.Tensor      # per-tile records (class-sorted pixel lists, counts)
    seg_label: torch.Tensor    # [N,P] uint8
    plane: int


def class_labels(logits: torch.Tensor, num_classes: int, channels: int, labels_val: Optional[torch.Tensor] = None,
                 overflow_channel: Optional[int] = None, conf_thresh: Optional[float] = None,
                 use_conf_gate: bool = False) -> ClassSums:
    """First launch of A5 alone: per-pixel argmax labels + tile records from the logits, and the [N,C,channels] sums
    zeroed by the same launch.  It needs no features, so a pipelined caller can run it for batch i+1 beside the
    kernels of batch i; ``class_accumulate`` completes the returned ClassSums."""
    x = _dev(logits, torch.float32, "logits")
    if x.dim() != 4 or x.shape[1] != num_classes:
        raise ValueError("logits must be [N,%d,h,w], got %s" % (num_classes, tuple(x.shape)))
    n, plane = x.shape[0], x[0, 0].numel()
    lv = None
    if labels_val is not None:
        lv = _dev(labels_val, torch.float32, "labels_val")
        if lv.numel() != n * plane:
            raise ValueError("labels_val must be [N,1,h,w]")
    dev = x.device
    L = lib()
    rec = torch.empty(max(int(L.tss_proto_records_bytes(n, plane)), 16), dtype=torch.uint8, device=dev)
    seg = torch.empty((n, plane), dtype=torch.uint8, device=dev)
    sums = torch.empty((n, num_classes, int(channels)), dtype=torch.float32, device=dev)
    check(L.tss_proto_labels(_ptr(x), _ptr(lv), n, num_classes, plane,
                             num_classes if overflow_channel is None else int(overflow_channel),
                             1 if use_conf_gate else 0, float(-1.0 if conf_thresh is None else conf_thresh),
                             _ptr(seg), _ptr(rec), _ptr(sums), sums.numel(), _stream()), "tss_proto_labels")
    return ClassSums(sums, rec, seg, plane)


def class_accumulate(cs: ClassSums, feat: torch.Tensor, consumer_warps: int = 0) -> ClassSums:
    """Second launch of A5 on the current stream: ONE streaming pass over feat into the (zeroed) sums of ``cs``."""
    f = _dev(feat, torch.float32, "feat")
    n, num_classes, k = cs.sums.shape
    if f.shape[0] != n or f.shape[1] != k or f[0, 0].numel() != cs.plane:
        raise ValueError("feat %s does not match the label records (N=%d, K=%d, plane=%d)" % (tuple(f.shape), n, k, cs.plane))
    check(lib().tss_proto_accumulate(_ptr(f), _ptr(cs.records), n, k, cs.plane, num_classes, _ptr(cs.sums), 0,
                                     int(consumer_warps), _stream()), "tss_proto_accumulate")
    return cs


def class_sums(feat: torch.Tensor, logits: torch.Tensor, num_classes: int, labels_val: Optional[torch.Tensor] = None,
               overflow_channel: Optional[int] = None, conf_thresh: Optional[float] = None, use_conf_gate: bool = False,
               consumer_warps: int = 0, accumulate_stream: Optional["torch.cuda.Stream"] = None) -> ClassSums:
    """First two launches of A5: per-pixel argmax labels + tile records, then ONE streaming pass over feat.
    ``consumer_warps`` 4 / 6 / 8 select the smaller forms of the accumulate kernel that share every SM with the
    pseudo-label kernel (``flags=pl_flags(...)``) running on another stream.  ``accumulate_stream``: the label kernel
    runs on the current stream, the accumulate kernel on that stream (ordered after it); work queued next on the
    current stream runs beside the accumulate kernel, and the caller joins
    (``current_stream().wait_stream(accumulate_stream)``) before anything reads the sums."""
    f = _dev(feat, torch.float32, "feat")
    x = _dev(logits, torch.float32, "logits")
    if x.dim() != 4 or f.dim() != 4 or x.shape[0] != f.shape[0] or x.shape[1] != num_classes or x[0, 0].numel() != f[0, 0].numel():
        raise ValueError("logits %s do not match feat %s / num_classes %d" % (tuple(x.shape), tuple(f.shape), num_classes))
    cs = class_labels(x, num_classes, f.shape[1], labels_val, overflow_channel, conf_thresh, use_conf_gate)
    if accumulate_stream is None:
        return class_accumulate(cs, f, consumer_warps)
    accumulate_stream.wait_stream(torch.cuda.current_stream(f.device))
    for t in (f, cs.records, cs.sums):
        t.record_stream(accumulate_stream)
    with torch.cuda.stream(accumulate_stream):
        return class_accumulate(cs, f, consumer_warps)


def class_finalize(cs: ClassSums, min_pixels: int = 10, update=None, scatter=None) -> ClassMeans:
    """Third launch of A5 on the current stream: means, validity rules, and either the sequential prototype update
    (``update=(protos, num, mode, start_mean, momentum[, cap])``, tss_proto_finalize_update) or the stores into every
    rank's gather buffer (``scatter=(vectors_out, sums_out, peer_ptrs, peer_sum_ptrs)``, tss_proto_finalize_scatter)."""
    sums, rec, plane = cs.sums, cs.records, cs.plane
    n, num_classes, k = sums.shape
    dev = sums.device
    L = lib()
    st = _stream()
    ca = torch.empty((n, num_classes), dtype=torch.int32, device=dev)
    cg = torch.empty((n, num_classes), dtype=torch.int32, device=dev)
    vec = torch.empty_like(sums)
    valid = torch.empty((n, num_classes), dtype=torch.uint8, device=dev)
    vsum = torch.empty((n, num_classes), dtype=torch.float32, device=dev)
    if scatter is not None:
        if update is not None:
            raise ValueError("scatter and update are exclusive (the update follows the exchange)")
        vec, vsum_out, peers, peer_sums = scatter
        if not (vec.is_cuda and vec.dtype == torch.float32 and vec.is_contiguous() and tuple(vec.shape) == (n, num_classes, k)):
            raise TssError("scatter: vectors_out must be a contiguous fp32 CUDA tensor [%d,%d,%d]" % (n, num_classes, k))
        if vsum_out is not None:
            if not (vsum_out.is_cuda and vsum_out.dtype == torch.float32 and vsum_out.is_contiguous() and vsum_out.numel() == n * num_classes):
                raise TssError("scatter: sums_out must be a contiguous fp32 CUDA tensor [%d,%d]" % (n, num_classes))
            vsum = vsum_out.view(n, num_classes)
        if peer_sums is not None and len(peer_sums) != len(peers):
            raise ValueError("scatter: peer_sum_ptrs must match peer_ptrs")
        arr = (ctypes.c_void_p * max(len(peers), 1))(*[ctypes.c_void_p(int(a)) for a in peers])
        arr_s = None if peer_sums is None else (ctypes.c_void_p * max(len(peers), 1))(*[ctypes.c_void_p(int(a)) for a in peer_sums])
        check(L.tss_proto_finalize_scatter(_ptr(sums), _ptr(rec), n, num_classes, k, plane, int(min_pixels), _ptr(ca),
                                           _ptr(cg), _ptr(vec), _ptr(valid), _ptr(vsum), arr, arr_s, len(peers), st),
              "tss_proto_finalize_scatter")
    elif update is None:
        check(L.tss_proto_finalize(_ptr(sums), _ptr(rec), n, num_classes, k, plane, int(min_pixels), _ptr(ca), _ptr(cg),
                                   _ptr(vec), _ptr(valid), _ptr(vsum), st), "tss_proto_finalize")
    else:
        protos, num, mode, start_mean, momentum = update[:5]
        cap = float(update[5]) if len(update) > 5 else 3000.0
        _check_bank(protos, num, num_classes, k)
        check(L.tss_proto_finalize_update(_ptr(sums), _ptr(rec), n, num_classes, k, plane, int(min_pixels), _ptr(ca),
                                          _ptr(cg), _ptr(vec), _ptr(valid), _ptr(vsum), _ptr(protos), _ptr(num), int(mode),
                                          1 if start_mean else 0, float(momentum), cap, st), "tss_proto_finalize_update")
    return ClassMeans(vec, valid, vsum, ca, cg, cs.seg_label, sums)


def class_means(feat: torch.Tensor, logits: torch.Tensor, num_classes: int, labels_val: Optional[torch.Tensor] = None,
                overflow_channel: Optional[int] = None, conf_thresh: Optional[float] = None,
                use_conf_gate: bool = False, min_pixels: int = 10, update=None, scatter=None,
                tail_stream: Optional["torch.cuda.Stream"] = None, consumer_warps: int = 0) -> ClassMeans:
    """A5 (three launches: labels -> segmented sums -> means) = class_sums + class_finalize.  ``update`` /
    ``scatter``: see class_finalize.  ``tail_stream``: issue the third launch (13 latency-bound CTAs at cfg2) on that
    stream, ordered after the first two, so that independent work queued next on the current stream runs beside it;
    the caller joins (``current_stream().wait_stream(tail_stream)``) before reading the means or the bank."""
    cs = class_sums(feat, logits, num_classes, labels_val, overflow_channel, conf_thresh, use_conf_gate, consumer_warps)
    if tail_stream is None:
        return class_finalize(cs, min_pixels, update, scatter)
    cur = torch.cuda.current_stream(cs.sums.device)
    tail_stream.wait_stream(cur)
    with torch.cuda.stream(tail_stream):
        out = class_finalize(cs, min_pixels, update, scatter)
    for t in (cs.records, cs.sums):                      # allocated under `cur`, read on the tail stream
        t.record_stream(tail_stream)
    for t in (out.vectors, out.valid, out.vector_sum, out.count_all, out.count_gated):
        if scatter is None or t is not out.vectors:      # allocated under the tail stream, read by the caller on `cur`
            t.record_stream(cur)
    return out


def prototype_update(protos: torch.Tensor, num: torch.Tensor, means: ClassMeans, mode: int, start_mean: bool,
                     momentum: float, cap: float = 3000.0) -> None:
    """In-place sequential update of protos [C,K] / num [C] with one batch (reference list order)."""
    for t, nm in ((protos, "protos"), (num, "num")):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise TssError("%s must be a contiguous fp32 CUDA tensor (updated in place)" % nm)
    n, c, k = means.vectors.shape
    if protos.shape != (c, k) or num.numel() != c:
        raise ValueError("protos %s / num %s do not match means [%d,%d,%d]" % (tuple(protos.shape), tuple(num.shape), n, c, k))
    check(lib().tss_proto_update(_ptr(protos), _ptr(num), _ptr(means.vectors), _ptr(means.valid), _ptr(means.vector_sum),
                                 n, c, k, mode, 1 if start_mean else 0, float(momentum), float(cap), _stream()),
          "tss_proto_update")


def prototype_update_partial(means: ClassMeans, num_before: torch.Tensor, mode: int, start_mean: bool, momentum: float,
                             first_index: Optional[torch.Tensor], total: Optional[torch.Tensor], cap: float = 3000.0):
    """Closed-form (sharded) update.  first_index=None: returns the per-class contribution counts [C] int32.
    Otherwise returns (partial [C,K], scale0 [C], num_after [C])."""
    n, c, k = means.vectors.shape
    dev = means.vectors.device
    L = lib()
    if first_index is None:
        cnt = torch.empty(c, dtype=torch.int32, device=dev)
        check(L.tss_proto_update_partial(_ptr(means.vectors), _ptr(means.valid), _ptr(means.vector_sum), n, c, k, mode,
                                         1 if start_mean else 0, float(momentum), float(cap), None, None, None, _ptr(cnt),
                                         None, None, None, _stream()), "tss_proto_update_partial")
        return cnt
    nb = _dev(num_before, torch.float32, "num_before")
    fi = _dev(first_index, torch.int32, "first_index")
    tot = _dev(total, torch.int32, "total")
    partial = torch.empty((c, k), dtype=torch.float32, device=dev)
    scale0 = torch.empty(c, dtype=torch.float32, device=dev)
    num_after = torch.empty(c, dtype=torch.float32, device=dev)
    check(L.tss_proto_update_partial(_ptr(means.vectors), _ptr(means.valid), _ptr(means.vector_sum), n, c, k, mode,
                                     1 if start_mean else 0, float(momentum), float(cap), _ptr(nb), _ptr(fi), _ptr(tot),
                                     None, _ptr(partial), _ptr(scale0), _ptr(num_after), _stream()),
          "tss_proto_update_partial")
    return partial, scale0, num_after


def prototype_update_single(protos: torch.Tensor, num: torch.Tensor, cls: int, vector: torch.Tensor, mode: int,
                            start_mean: bool, momentum: float, cap: float = 3000.0) -> None:
    """update_objective_SingleVector(id, vector): one vector into row `cls`, in place, no host sync (the
    `vector.sum() == 0` skip is evaluated inside the kernel)."""
    for t, nm in ((protos, "protos"), (num, "num")):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise TssError("%s must be a contiguous fp32 CUDA tensor (updated in place)" % nm)
    c, k = protos.shape
    if not 0 <= int(cls) < c:
        raise IndexError("class id %d outside [0, %d)" % (cls, c))
    v = _dev(vector, torch.float32, "vector").reshape(-1)
    if v.numel() != k:
        raise ValueError("vector has %d elements, prototypes have %d channels" % (v.numel(), k))
    row = ctypes.c_void_p(protos.data_ptr() + int(cls) * k * 4)
    cnt = ctypes.c_void_p(num.data_ptr() + int(cls) * 4)
    check(lib().tss_proto_update(row, cnt, _ptr(v), None, None, 1, 1, k, mode, 1 if start_mean else 0,
                                 float(momentum), float(cap), _stream()), "tss_proto_update")


def one_hot_overflow(label: torch.Tensor, num_classes: int, overflow_channel: int) -> torch.Tensor:
    """process_label: float label map [N,1,h,w] -> one-hot [N,C+1,h,w] fp32."""
    lab = _dev(label, torch.float32, "label")
    if lab.dim() != 4 or lab.shape[1] != 1:
        raise ValueError("label must be [N,1,h,w]")
    n, _, a, b = lab.shape
    out = torch.empty((n, num_classes + 1, a, b), dtype=torch.float32, device=lab.device)
    check(lib().tss_one_hot_overflow(_ptr(lab), n, a * b, num_classes, int(overflow_channel), _ptr(out), _stream()),
          "tss_one_hot_overflow")
    return out


def weak_param_table(weak_params, n: int):
    """The reference's int(./4) arithmetic on weak_params (self_training.py:148-152) -> rows
    {rh, rw, y0, x0, flip, 0} plus each image's crop extent after Python's slice clipping."""
    sized, crop, flip = weak_params["RandomSized"], weak_params["RandomCrop"], weak_params["RandomHorizontallyFlip"]
    rows, extents = [], []
    for i in range(n):
        hh, ww = float(sized[0][i]), float(sized[1][i])
        rh, rw = int(hh / 4), int(ww / 4)
        y1, y2, x1, x2 = (float(crop[j][i]) for j in range(4))
        ys, th, xs, tw = int(y1 / 4), int((y2 - y1) / 4), int(x1 / 4), int((x2 - x1) / 4)
        th_eff = max(0, min(ys + th, rh) - min(ys, rh))
        tw_eff = max(0, min(xs + tw, rw) - min(xs, rw))
        rows.append([rh, rw, ys, xs, 1 if bool(flip[i]) else 0, 0])
        extents.append((th_eff, tw_eff))
    return np.asarray(rows, np.int32).reshape(n, 6), extents


@dataclass
class WeakTable:
    """weak_params of one batch after the reference's int(./4) arithmetic, resident on the device: build it once per
    batch (weak_table) and hand it to full2weak / prototype_distance_sampled -- e.g. outside a CUDA-graph capture."""
    params: torch.Tensor       # [N,6] int32 on device
    th: int
    tw: int


def weak_table(weak_params, n: int, device) -> WeakTable:
    if isinstance(weak_params, WeakTable):
        return weak_params
    table, extents = weak_param_table(weak_params, n)
    if len(set(extents)) != 1:
        raise RuntimeError("full2weak: crops differ in size across the batch %s (torch.cat raises in the reference)"
                           % sorted(set(extents)))
    th, tw = extents[0]
    if th <= 0 or tw <= 0:
        raise RuntimeError("full2weak: empty crop %s" % (extents[0],))
    if (table[:, 0] < 1).any() or (table[:, 1] < 1).any():
        raise RuntimeError("full2weak: resized extent below 1")
    return WeakTable(torch.from_numpy(table).to(device), th, tw)


def full2weak(feat: torch.Tensor, weak_params) -> torch.Tensor:
    """SelfTrain.full2weak in one gather pass (tss_full2weak).  ``weak_params``: the reference's dict or a WeakTable."""
    f = _dev(feat, torch.float32, "feat")
    n, k, h, w = f.shape
    wt = weak_table(weak_params, n, f.device)
    params, th, tw = wt.params, wt.th, wt.tw
    out = torch.empty((n, k, th, tw), dtype=torch.float32, device=f.device)
    check(lib().tss_full2weak(_ptr(f), n, k, h, w, _ptr(params), th, tw, _ptr(out), _stream()), "tss_full2weak")
    return out


# ------------------------------------------------------------------------------------------- A7/A8/A10
def prototype_distance(feat: torch.Tensor, protos: torch.Tensor, temperature: float = 1.0, want_dist: bool = True,
                       want_weight: bool = False, soft: Optional[torch.Tensor] = None,
                       hard: Optional[torch.Tensor] = None, train_thred: float = 0.0, ignore_index: int = 255,
                       form: int = 0):
    """Returns (dist [N,C,h,w] | None, weight [N,C,h,w] | None, label [N,1,h,w] int64 | None)."""
    f = _dev(feat, torch.float32, "feat")
    p = _dev(protos, torch.float32, "protos")
    n, k, h, w = f.shape
    c = p.shape[0]
    if p.shape[1] != k:
        raise ValueError("protos [C,%d] do not match feat channels %d" % (p.shape[1], k))
    plane = h * w
    dev = f.device
    dist = torch.empty((n, c, h, w), dtype=torch.float32, device=dev) if want_dist else None
    wt = torch.empty((n, c, h, w), dtype=torch.float32, device=dev) if want_weight else None
    sf = hd = lab = None
    hd_dt = 0
    if soft is not None:
        sf = _dev(soft, torch.float32, "soft")
        if sf.shape != (n, c, h, w):
            raise ValueError("soft %s must be [N,C,h,w] = %s" % (tuple(sf.shape), (n, c, h, w)))
        lab = torch.empty((n, 1, h, w), dtype=torch.int64, device=dev)
        if hard is not None:
            hd = _label_tensor(hard, "hard")
            if hd.numel() != n * plane:
                raise ValueError("hard must have N*h*w elements")
            hd_dt = _DT[hd.dtype]
    check(lib().tss_proto_distance(_ptr(f), _ptr(p), n, k, plane, c, float(temperature), _ptr(dist), _ptr(wt), _ptr(sf),
                                   _ptr(hd), hd_dt, float(train_thred), int(ignore_index), _ptr(lab), int(form), _stream()),
          "tss_proto_distance")
    return dist, wt, lab


def prototype_distance_sampled(feat8: torch.Tensor, quarter_hw: Sequence[int], weak_params, protos: torch.Tensor,
                               temperature: float = 1.0, want_dist: bool = False, want_weight: bool = False,
                               soft: Optional[torch.Tensor] = None, hard: Optional[torch.Tensor] = None,
                               train_thred: float = 0.0, ignore_index: int = 255, flags: int = 0):
    """get_prototype_weight / the rectify block on the network's own 1/8-stride features: the interpolation to the
    quarter map (self_training.py:69) and full2weak (:145-158) are evaluated inside the distance kernel
    (tss_proto_distance_sampled).  Returns (dist, weight, label) at the weak view's [N, ., TH, TW]."""
    f = _dev(feat8, torch.float32, "feat8")
    p = _dev(protos, torch.float32, "protos")
    n, k, h8, w8 = f.shape
    c = p.shape[0]
    if p.shape[1] != k:
        raise ValueError("protos [C,%d] do not match feat channels %d" % (p.shape[1], k))
    Hq, Wq = int(quarter_hw[0]), int(quarter_hw[1])
    dev = f.device
    wtab = weak_table(weak_params, n, dev)
    params, th, tw = wtab.params, wtab.th, wtab.tw
    dist = torch.empty((n, c, th, tw), dtype=torch.float32, device=dev) if want_dist else None
    wt = torch.empty((n, c, th, tw), dtype=torch.float32, device=dev) if want_weight else None
    sf = hd = lab = None
    hd_dt = 0
    if soft is not None:
        sf = _dev(soft, torch.float32, "soft")
        if sf.shape != (n, c, th, tw):
            raise ValueError("soft %s must be [N,C,h,w] = %s" % (tuple(sf.shape), (n, c, th, tw)))
        lab = torch.empty((n, 1, th, tw), dtype=torch.int64, device=dev)
        if hard is not None:
            hd = _label_tensor(hard, "hard")
            if hd.numel() != n * th * tw:
                raise ValueError("hard must have N*h*w elements")
            hd_dt = _DT[hd.dtype]
    check(lib().tss_proto_distance_sampled(_ptr(f), _ptr(p), n, k, h8, w8, Hq, Wq, _ptr(params), th, tw, c, float(temperature),
                                           _ptr(dist), _ptr(wt), _ptr(sf), _ptr(hd), hd_dt, float(train_thred),
                                           int(ignore_index), _ptr(lab), int(flags), _stream()), "tss_proto_distance_sampled")
    return dist, wt, lab


FUSED_SAMPLING_MAX_WORK = 1 << 24   # pixel-channels: above this the two materialising passes + streaming distance kernel win


def rectify_from_stride8(feat8: torch.Tensor, quarter_hw: Sequence[int], weak_params, protos: torch.Tensor,
                         temperature: float, soft: torch.Tensor, hard: Optional[torch.Tensor], train_thred: float,
                         ignore_index: int, want_weight: bool = False, fused: Optional[bool] = None):
    """self_training.py:69 + :82 + :103-112 from the network's 1/8-stride features: (weights | None, labels).
    Small problems (the reference's own batches: 2-4 frames, 64x128 crops, 256 channels) run the ONE fused sampling
    kernel -- no intermediate tensor; large batches (measured on B200: 16 x 256 x 80 x 160 takes 688 us fused against
    119 + 74 us) interpolate to the quarter map, apply full2weak and go through the streaming distance kernel, whose
    feature loads are TMA bulk copies.  Same arithmetic either way (both reproduce the two ATen interpolation stages)."""
    n, k = feat8.shape[0], feat8.shape[1]
    wt = weak_table(weak_params, n, feat8.device)
    if fused is None:
        fused = n * k * wt.th * wt.tw <= FUSED_SAMPLING_MAX_WORK
    if fused:
        _, w, lab = prototype_distance_sampled(feat8, quarter_hw, wt, protos, temperature=temperature, want_weight=want_weight,
                                               soft=soft, hard=hard, train_thred=train_thred, ignore_index=ignore_index)
        return w, lab
    fw = full2weak(upsample_bilinear(feat8, quarter_hw), wt)
    _, w, lab = prototype_distance(fw, protos, temperature=temperature, want_dist=False, want_weight=want_weight, soft=soft,
                                   hard=hard, train_thred=train_thred, ignore_index=ignore_index)
    return w, lab


# ------------------------------------------------------------------------------------------- A11-A13
_LOSS_WS = {}


def _loss_workspace(dev) -> torch.Tensor:
    """Partials + ticket counter of the loss kernels.  The ticket must be zero at launch and every launch leaves it zero:
    one zero-initialised workspace per (device, stream) is allocated once and reused.  Inside a CUDA-graph capture a
    fresh zeroed workspace is taken from the graph's pool (the zero fill is a node of the graph)."""
    nbytes = int(lib().tss_loss_workspace_bytes())
    if torch.cuda.is_current_stream_capturing():
        return torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    key = (torch.device(dev).index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _LOSS_WS.get(key)
    if ws is None:
        ws = _LOSS_WS[key] = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    return ws


class _CeRce(torch.autograd.Function):
    """alpha * CE(ignore_index, mean) + beta * RCE, forward and input gradient in the CUDA kernels."""

    @staticmethod
    def forward(ctx, logits, target, ignore_index, alpha, beta, with_rce, ce_sum=False):
        x = _dev(logits, torch.float32, "logits")
        t = _label_tensor(target, "target")
        n, c = x.shape[0], x.shape[1]
        plane = x[0, 0].numel()
        if t.numel() != n * plane:
            raise ValueError("target must have N*H*W elements")
        out = torch.empty(8, dtype=torch.float64, device=x.device)
        losses = torch.empty(3, dtype=torch.float32, device=x.device)
        # ce = sum / valid (nan when nothing is valid, as torch), rce = sum / (valid + 1e-6), total = alpha ce + beta rce:
        # formed inside the reduction launch (tss_ce_rce_fwd_mix) instead of eight scalar torch kernels
        check(lib().tss_ce_rce_fwd_mix(_ptr(x), _ptr(t), _DT[t.dtype], n, c, plane, int(ignore_index), 1 if with_rce else 0,
                                       float(alpha), float(beta), 1 if ce_sum else 0, _ptr(out), _ptr(losses),
                                       _ptr(_loss_workspace(x.device)), _stream()), "tss_ce_rce_fwd_mix")
        ctx.save_for_backward(x, t, out)
        ctx.meta = (int(ignore_index), float(alpha), float(beta), bool(with_rce), bool(ce_sum))
        return losses[0], losses[1], losses[2], out[3], out[4]

    @staticmethod
    def backward(ctx, g_total, g_ce, g_rce, _g_bad, _g_over):
        x, t, out = ctx.saved_tensors
        ignore_index, alpha, beta, with_rce, ce_sum = ctx.meta
        n, c = x.shape[0], x.shape[1]
        plane = x[0, 0].numel()
        # ce_scale = (g_total alpha + g_ce) / valid, rce_scale = (g_total beta + g_rce) / (valid + 1e-6): formed in the kernel
        up = torch.stack([g_total, g_ce, g_rce]).to(torch.float32)
        grad = torch.empty_like(x)
        check(lib().tss_ce_rce_bwd_mix(_ptr(x), _ptr(t), _DT[t.dtype], n, c, plane, ignore_index, 1 if with_rce else 0,
                                       float(alpha), float(beta), 1 if ce_sum else 0, _ptr(up), _ptr(out), _ptr(grad),
                                       _stream()), "tss_ce_rce_bwd_mix")
        return grad, None, None, None, None, None, None


def cross_entropy(logits, target, ignore_index: int = -100, reduction: str = "mean") -> torch.Tensor:
    """nn.CrossEntropyLoss(ignore_index, reduction='mean'|'sum') on [N,C,...] logits."""
    if reduction not in ("mean", "sum"):
        raise ValueError("reduction must be 'mean' or 'sum'")
    return _CeRce.apply(logits, target, ignore_index, 1.0, 0.0, False, reduction == "sum")[0]


def cross_entropy_checked(logits, target, ignore_index: int = -100, reduction: str = "mean"):
    """(loss, bad_label_count): the count (device scalar) of targets outside [0, C) other than ignore_index, which
    torch's CrossEntropyLoss refuses (IndexError on CPU, device assert on CUDA)."""
    out = _CeRce.apply(logits, target, ignore_index, 1.0, 0.0, False, reduction == "sum")
    return out[0], out[3]


def ce_rce(logits, target, ignore_index: int, alpha: float, beta: float):
    """(alpha*CE + beta*RCE, CE, RCE, bad_label_count, label_equals_C_count) -- self_training.py:114-118.  A
    label equal to C passes SelfTrain.rce's one_hot(label, C+1) (it adds -log(1e-4) sum_c p_c and counts in the
    mask) but is out of bounds for the CE term; it is reported in the last output, never folded into CE."""
    return _CeRce.apply(logits, target, ignore_index, alpha, beta, True)


def reverse_cross_entropy(logits, target, ignore_index: int) -> torch.Tensor:
    return _CeRce.apply(logits, target, ignore_index, 0.0, 1.0, True)[2]


class _Focal(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, gamma, alpha, size_average):
        x = _dev(logits, torch.float32, "logits")
        t = _label_tensor(target, "target")
        n, c = x.shape[0], x.shape[1]
        plane = x[0, 0].numel() if x.dim() > 2 else 1
        if x.dim() == 2:  # [M, C] rows: treat as N=M images of one pixel
            plane = 1
        a = None if alpha is None else _dev(alpha, torch.float32, "alpha")
        out = torch.empty(8, dtype=torch.float64, device=x.device)
        check(lib().tss_focal_fwd(_ptr(x), _ptr(t), _DT[t.dtype], n, c, plane, float(gamma), _ptr(a), _ptr(out),
                                  _ptr(_loss_workspace(x.device)), _stream()), "tss_focal_fwd")
        ctx.save_for_backward(x, t, out, a if a is not None else torch.empty(0, device=x.device))
        ctx.meta = (float(gamma), a is not None, bool(size_average))
        loss = out[0] / out[1] if size_average else out[0]
        return loss.to(torch.float32), out[3]

    @staticmethod
    def backward(ctx, g, _g_bad):
        x, t, out, a = ctx.saved_tensors
        gamma, has_alpha, size_average = ctx.meta
        n, c = x.shape[0], x.shape[1]
        plane = x[0, 0].numel() if x.dim() > 2 else 1
        scale = (g.double() / out[1] if size_average else g.double()).to(torch.float32).reshape(1).contiguous()
        grad = torch.empty_like(x)
        check(lib().tss_focal_bwd(_ptr(x), _ptr(t), _DT[t.dtype], n, c, plane, gamma, _ptr(a) if has_alpha else None,
                                  _ptr(scale), _ptr(grad), _stream()), "tss_focal_bwd")
        return grad, None, None, None, None


def focal_loss(logits, target, gamma: float = 0.0, alpha: Optional[torch.Tensor] = None, size_average: bool = True):
    return _Focal.apply(logits, target, gamma, alpha, size_average)[0]


def launch_count() -> int:
    return int(lib().tss_launch_count())
