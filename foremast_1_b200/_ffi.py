"""ctypes binding of libforemast_cuda.so — the same C ABI the cgo package binds.

The library is the only compute path of this package: if it is missing, or no
CUDA device is present, every scoring call raises (there is no CPU fallback).
"""
import ctypes as C
import os

from . import _abi as A
from . import build as _build

_lib = None

# name -> (restype, argtypes); every symbol declared in include/foremast_cuda.h
SYMBOLS = {
    "fm_version": (C.c_int32, []),
    "fm_sizeof": (C.c_uint32, [C.c_uint32]),
    "fm_device_count": (C.c_int32, [C.POINTER(C.c_int32)]),
    "fm_ctx_create": (C.c_int32, [C.c_int32, C.POINTER(C.c_void_p)]),
    "fm_ctx_destroy": (C.c_int32, [C.c_void_p]),
    "fm_default_params": (C.c_int32, [C.POINTER(A.GlobalParams), C.POINTER(A.TypeParams), C.POINTER(C.c_uint32)]),
    "fm_set_params": (C.c_int32, [C.c_void_p, C.POINTER(A.GlobalParams), C.POINTER(A.TypeParams), C.c_uint32]),
    "fm_alloc_pinned": (C.c_int32, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "fm_free_pinned": (C.c_int32, [C.c_void_p]),
    "fm_device_alloc": (C.c_int32, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "fm_device_free": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "fm_memcpy_h2d": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "fm_memcpy_d2h": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "fm_score_batch": (C.c_int32, [C.c_void_p, C.POINTER(A.Batch), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(A.Timing)]),
    "fm_score_batch_cached": (C.c_int32, [C.c_void_p, C.POINTER(A.Batch), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(A.Timing)]),
    "fm_cache_configure": (C.c_int32, [C.c_void_p, C.c_uint64]),
    "fm_cache_query": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "fm_cache_get_stats": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "fm_score_pairs": (C.c_int32, [C.c_void_p, C.POINTER(A.PairBatch), C.c_void_p, C.POINTER(A.Timing)]),
    "fm_score_pairs_device": (C.c_int32, [C.c_void_p, C.POINTER(A.PairBatch), C.c_void_p, C.POINTER(A.Timing)]),
    "fm_pool_create": (C.c_int32, [C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_void_p)]),
    "fm_pool_destroy": (C.c_int32, [C.c_void_p]),
    "fm_pool_size": (C.c_int32, [C.c_void_p]),
    "fm_pool_ctx": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "fm_pool_set_params": (C.c_int32, [C.c_void_p, C.POINTER(A.GlobalParams), C.POINTER(A.TypeParams), C.c_uint32]),
    "fm_pool_cache_configure": (C.c_int32, [C.c_void_p, C.c_uint64]),
    "fm_pool_score_batch": (C.c_int32, [C.c_void_p, C.POINTER(A.Batch), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fm_pool_last_error": (C.c_char_p, [C.c_void_p]),
    "fm_score_batch_device": (C.c_int32, [C.c_void_p, C.POINTER(A.Batch), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(A.Timing)]),
    "fm_score_groups": (C.c_int32, [C.c_void_p, C.POINTER(A.GroupBatch), C.c_void_p, C.c_void_p, C.POINTER(A.Timing)]),
    "fm_score_groups_device": (C.c_int32, [C.c_void_p, C.POINTER(A.GroupBatch), C.c_void_p, C.c_void_p, C.POINTER(A.Timing)]),
    "fm_rank_dump": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p]),
    "fm_ma_window_series": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
    "fm_ma_window_series_static": (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
    "fm_synth_fill_device": (C.c_int32, [C.c_void_p, C.POINTER(A.SynthDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fm_decode_query_range": (C.c_int64, [C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                          C.c_uint32, C.POINTER(C.c_uint32)]),
    "fm_synth_fill_host": (C.c_int32, [C.POINTER(A.SynthDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "fm_ctx_sync": (C.c_int32, [C.c_void_p]),
    "fm_last_error": (C.c_char_p, [C.c_void_p]),
}


class ForemastCudaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libforemast_cuda error {code}: {msg}")
        self.code = code


def lib_path():
    return _build.SO


def lib():
    """Load (building first if the sources are newer) libforemast_cuda.so."""
    global _lib
    if _lib is None:
        path = os.environ.get("FOREMAST_CUDA_LIB") or _build.SO   # override: an already built library (kernel variants)
        if path == _build.SO and _build.stale():
            try:
                _build.build()
            except Exception as e:  # no nvcc on this machine: use the shipped .so if any
                if not os.path.exists(path):
                    raise RuntimeError(f"libforemast_cuda.so is not built and cannot be built here: {e}")
        _lib = C.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(rc, ctx=None):
    if rc != A.FM_OK:
        msg = lib().fm_last_error(ctx)
        raise ForemastCudaError(rc, (msg or b"").decode("utf-8", "replace"))
