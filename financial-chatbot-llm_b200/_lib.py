"""ctypes binding of libfcb200.so — the C-ABI declared in include/fcb200.h.

This is the binding a maintainer of the reference adds in place of
``from qdrant_client import QdrantClient as QdrantHttpClient``
(reference tools/qdrant_tool.py:2); see INTEGRATION.md. There is no CPU path:
loading fails loudly when the library has not been built
(``python -c "import __graft_entry__ as g; g.build()"``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FCB_LIB_PATH") or os.path.join(_HERE, "csrc", "libfcb200.so")   # override: A/B of two builds

FCB_OK = 0
FCB_ERR_INVALID, FCB_ERR_CUDA, FCB_ERR_NOMEM, FCB_ERR_STATE, FCB_ERR_UNSUPPORTED = -1, -2, -3, -4, -5
FCB_DTYPE_F32, FCB_DTYPE_BF16 = 0, 1
FCB_MAX_K = 10000
FCB_DATE_MIN, FCB_DATE_MAX = -(1 << 63), (1 << 63) - 1
FCB_ALL_CATEGORIES = (1 << 64) - 1
ABI_VERSION = 1


class FcbConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_batch", C.c_int32), ("max_k", C.c_int32),
                ("reserved", C.c_int32)]


class FcbShard(C.Structure):
    _fields_ = [("vectors", C.c_void_p), ("ids", C.c_void_p), ("id_rank", C.c_void_p),
                ("dates", C.c_void_p), ("categories", C.c_void_p), ("cents", C.c_void_p),
                ("seg_offsets", C.c_void_p),
                ("n_rows", C.c_int64), ("n_segs", C.c_int64), ("row_base", C.c_int64),
                ("dim", C.c_int32), ("dtype", C.c_int32)]


class FcbFilter(C.Structure):
    _fields_ = [("row_begin", C.c_int64), ("row_end", C.c_int64),
                ("date_lo", C.c_int64), ("date_hi", C.c_int64),
                ("category_mask", C.c_uint64)]


FCB_MAX_WORLD = 8
FCB_PEER_HANDLE_BYTES = 64


class FcbExchange(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("bufs", C.c_void_p * FCB_MAX_WORLD), ("cuts", C.c_void_p)]


class FcbError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"fcb200 error {code}: {message}")
        self.code = code


# name -> (restype, argtypes); every symbol include/fcb200.h declares
SIGNATURES = {
    "fcb_abi_version": (C.c_int32, []),
    "fcb_open": (C.c_int32, [C.POINTER(FcbConfig), C.POINTER(C.c_void_p)]),
    "fcb_close": (C.c_int32, [C.c_void_p]),
    "fcb_last_error": (C.c_char_p, [C.c_void_p]),
    "fcb_load_shard": (C.c_int32, [C.c_void_p, C.POINTER(FcbShard)]),
    "fcb_search": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(FcbFilter), C.c_int32, C.c_int32,
                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fcb_search_host": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(FcbFilter), C.c_int32,
                                    C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fcb_search_noalloc": (C.c_int32, [C.c_int32, C.POINTER(FcbShard), C.c_void_p, C.POINTER(FcbFilter),
                                       C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_char_p, C.c_int32]),
    "fcb_last_search_ms": (C.c_float, [C.c_void_p]),
    "fcb_last_search_launches": (C.c_int32, [C.c_void_p]),
    "fcb_scan_config": (C.c_int32, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "fcb_debug_scan_timing": (C.c_int32, [C.c_void_p]),
    "fcb_debug_dense_prof": (C.c_int32, [C.c_void_p]),
    "fcb_debug_dense_dump": (C.c_int32, [C.c_void_p, C.c_int64]),
    "fcb_last_dense_redo": (C.c_int32, [C.c_void_p, C.c_int32]),
    "fcb_last_dense_widened": (C.c_int32, [C.c_void_p, C.c_int32]),
    "fcb_exact_reruns": (C.c_int64, [C.c_void_p]),
    "fcb_exchange_bytes": (C.c_int64, [C.c_int32, C.c_int32]),
    "fcb_peer_alloc": (C.c_int32, [C.c_int32, C.c_int64, C.POINTER(C.c_void_p), C.c_void_p]),
    "fcb_peer_open": (C.c_int32, [C.c_int32, C.c_void_p, C.POINTER(C.c_void_p)]),
    "fcb_peer_close": (C.c_int32, [C.c_int32, C.c_void_p, C.c_int32]),
    "fcb_debug_exchange_phases": (C.c_int32, [C.c_int32]),
    "fcb_search_routed": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_uint32,
                                      C.c_void_p, C.c_void_p]),
    "fcb_aggregate": (C.c_int32, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fcb_aggregate_columns": (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                          C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
    "fcb_group_keys_build": (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_int32),
                                         C.c_void_p]),
    "fcb_aggregate_columns_keyed": (C.c_int32, [C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int64, C.c_int64, C.c_int32,
                                                C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fcb_aggregate_hits": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                       C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fcb_merge_topk": (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                   C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fcb_shard_dates_sorted": (C.c_int32, [C.c_void_p]),
    "fcb_synth_dates_sorted": (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                           C.c_int64, C.c_void_p]),
    "fcb_synth_vectors": (C.c_int32, [C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_int64,
                                      C.c_int32, C.c_uint64, C.c_int32, C.c_void_p]),
    "fcb_synth_columns": (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_int64, C.c_uint64, C.c_int32, C.c_int64, C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the library and bind every declared symbol. Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FcbError(FCB_ERR_STATE,
                           f"{LIB_PATH} not built; run __graft_entry__.build() (nvcc, sm_100a). "
                           "There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        if lib.fcb_abi_version() != ABI_VERSION:
            raise FcbError(FCB_ERR_STATE, "libfcb200.so ABI version mismatch; rebuild")
        _lib = lib
    return _lib


def check(rc: int, handle=None) -> None:
    if rc != FCB_OK:
        msg = load().fcb_last_error(handle)
        raise FcbError(rc, msg.decode() if msg else "unknown")
