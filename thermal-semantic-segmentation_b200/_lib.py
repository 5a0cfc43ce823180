"""ctypes binding of libtss_b200.so (include/tss_b200.h).  There is NO fallback: if the shared object
is missing and cannot be built with nvcc, importing the ops fails loudly."""
from __future__ import annotations

import ctypes as C
import os

from . import _build

c_i64, c_int, c_f32, c_f64, c_vp, c_sz = C.c_int64, C.c_int, C.c_float, C.c_double, C.c_void_p, C.c_size_t


class TssError(RuntimeError):
    """A non-zero status from the C ABI (argument rejection or CUDA failure)."""


_SIGNATURES = {
    # name: (restype, [argtypes])
    "tss_version": (c_int, []),
    "tss_last_error_string": (C.c_char_p, []),
    "tss_source_digest": (C.c_char_p, []),
    "tss_launch_count": (C.c_uint64, []),
    "tss_confusion": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_vp, c_vp, c_int, c_vp]),
    "tss_softmax_nchw": (c_int, [c_vp, c_vp, c_i64, c_int, c_i64, c_vp]),
    "tss_pseudo_label": (c_int, [c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp,
                                 c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_vp]),
    "tss_pseudo_label_hist": (c_int, [c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp,
                                      c_vp, c_vp, c_vp, c_vp, c_int, c_vp, c_vp, c_int, c_vp]),
    "tss_upsample_bilinear": (c_int, [c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_vp]),
    "tss_upsample_bilinear_adjoint": (c_int, [c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_vp]),
    "tss_threshold_labels": (c_int, [c_vp, c_vp, c_i64, c_int, c_vp, c_vp, c_int, c_vp, c_vp]),
    "tss_gather_nearest_u8": (c_int, [c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp]),
    "tss_proto_records_bytes": (c_sz, [c_i64, c_i64]),
    "tss_proto_labels": (c_int, [c_vp, c_vp, c_i64, c_int, c_i64, c_int, c_int, c_f32, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "tss_proto_accumulate": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_vp, c_int, c_int, c_vp]),
    "tss_proto_finalize": (c_int, [c_vp, c_vp, c_i64, c_int, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tss_class_means_noalloc": (c_int, [c_vp, c_vp, c_i64, c_int, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tss_proto_finalize_scatter": (c_int, [c_vp, c_vp, c_i64, c_int, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_int, c_vp]),
    "tss_proto_finalize_update": (c_int, [c_vp, c_vp, c_i64, c_int, c_i64, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp,
                                          c_vp, c_vp, c_int, c_int, c_f64, c_f64, c_vp]),
    "tss_proto_update": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_i64, c_int, c_int, c_f64, c_f64, c_vp]),
    "tss_proto_update_partial": (c_int, [c_vp, c_vp, c_vp, c_i64, c_int, c_i64, c_int, c_int, c_f64, c_f64, c_vp,
                                         c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tss_proto_distance": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_f32, c_vp, c_vp, c_vp, c_vp, c_int,
                                   c_f32, c_int, c_vp, c_int, c_vp]),
    "tss_proto_distance_sampled": (c_int, [c_vp, c_vp, c_i64, c_i64, c_int, c_int, c_int, c_int, c_vp, c_int, c_int, c_int,
                                           c_f32, c_vp, c_vp, c_vp, c_vp, c_int, c_f32, c_int, c_vp, c_int, c_vp]),
    "tss_full2weak": (c_int, [c_vp, c_i64, c_i64, c_int, c_int, c_vp, c_int, c_int, c_vp, c_vp]),
    "tss_one_hot_overflow": (c_int, [c_vp, c_i64, c_i64, c_int, c_int, c_vp, c_vp]),
    "tss_loss_workspace_bytes": (c_sz, []),
    "tss_ce_rce_fwd": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_i64, c_int, c_int, c_vp, c_vp, c_vp]),
    "tss_ce_rce_bwd_mix": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_i64, c_int, c_int, c_f64, c_f64, c_int, c_vp, c_vp, c_vp, c_vp]),
    "tss_ce_rce_fwd_mix": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_i64, c_int, c_int, c_f64, c_f64, c_int, c_vp, c_vp, c_vp, c_vp]),
    "tss_ce_rce_bwd": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_i64, c_int, c_vp, c_vp, c_vp]),
    "tss_focal_fwd": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_i64, c_f32, c_vp, c_vp, c_vp, c_vp]),
    "tss_focal_bwd": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_i64, c_f32, c_vp, c_vp, c_vp, c_vp]),
}

EXPORTED = tuple(_SIGNATURES)

_lib = None


def library_path() -> str:
    return _build.LIB_PATH


def lib() -> C.CDLL:
    """The loaded library.  It is (re)built with nvcc when it is missing or was compiled from other sources than
    the ones in csrc/ (the source digest is compiled into the library); without nvcc a stale or missing library
    is an error -- kernels never run against old code silently, and there is no other code path to fall back to."""
    global _lib
    if _lib is None:
        try:
            path = _build.ensure_current()
        except RuntimeError as e:
            raise TssError("libtss_b200.so is missing or stale and cannot be rebuilt: %s" % e) from e
        handle = C.CDLL(path)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so does not export what the header declares
            fn.restype, fn.argtypes = res, args
        if handle.tss_version() != 100:
            raise TssError("libtss_b200.so version mismatch: %d" % handle.tss_version())
        _lib = handle
    return _lib


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = lib().tss_last_error_string()
        raise TssError("%s failed with status %d: %s" % (what or "libtss_b200", status, (msg or b"").decode()))
