"""ORACLE — test infrastructure only. ctypes binding of oracle/canon.c (see its header)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libcanon.so")


class CanonFilter(C.Structure):
    _fields_ = [("row_begin", C.c_int64), ("row_end", C.c_int64),
                ("date_lo", C.c_int64), ("date_hi", C.c_int64),
                ("category_mask", C.c_uint64)]


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "canon.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "libcanon.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.canon_scores.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
        L.canon_scores.restype = None
        L.canon_search.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_uint64,
                                   C.c_void_p, C.c_void_p, C.POINTER(CanonFilter), C.c_void_p,
                                   C.c_int64, C.c_void_p, C.c_void_p]
        L.canon_search.restype = C.c_int64
        L.canon_month_index.argtypes = [C.c_int64]
        L.canon_month_index.restype = C.c_int32
        L.canon_groupby.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p]
        L.canon_groupby.restype = C.c_int64
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _vec(vectors):
    """float32 [n,d] -> dtype 0; uint16 [n,d] (bf16 bit patterns) -> dtype 1."""
    v = np.ascontiguousarray(vectors)
    if v.dtype == np.float32:
        return v, 0
    if v.dtype == np.uint16:
        return v, 1
    raise TypeError("vectors must be float32 or uint16 (bf16 bits)")


def scores(vectors, query) -> np.ndarray:
    v, dt = _vec(vectors)
    q = np.ascontiguousarray(query, dtype=np.float64)
    out = np.empty(v.shape[0], dtype=np.float64)
    lib().canon_scores(_ptr(v), dt, v.shape[0], v.shape[1], _ptr(q), _ptr(out))
    return out


ALL_DATES = (-(1 << 62), 1 << 62)
ALL_CATEGORIES = (1 << 64) - 1


def search(vectors, query, k, *, row_begin=0, row_end=None, ids=None, id_base=0, dates=None,
           categories=None, date_lo=None, date_hi=None, category_mask=None):
    """Filtered exact top-k. Returns (ids uint64[m], scores float64[m]), m <= k."""
    v, dt = _vec(vectors)
    n, d = v.shape
    q = np.ascontiguousarray(query, dtype=np.float64)
    f = CanonFilter(row_begin, n if row_end is None else row_end,
                    ALL_DATES[0] if date_lo is None else int(date_lo),
                    ALL_DATES[1] if date_hi is None else int(date_hi),
                    ALL_CATEGORIES if category_mask is None else int(category_mask))
    ids_a = None if ids is None else np.ascontiguousarray(ids, dtype=np.uint64)
    dates_a = None if dates is None else np.ascontiguousarray(dates, dtype=np.int32)
    cats_a = None if categories is None else np.ascontiguousarray(categories, dtype=np.uint8)
    k = int(k)
    oi = np.empty(max(k, 1), dtype=np.uint64)
    os_ = np.empty(max(k, 1), dtype=np.float64)
    m = lib().canon_search(_ptr(v), dt, n, d, _ptr(ids_a), id_base, _ptr(dates_a), _ptr(cats_a),
                           C.byref(f), _ptr(q), k, _ptr(oi), _ptr(os_))
    if m < 0:
        raise MemoryError("canon_search")
    return oi[:m].copy(), os_[:m].copy()


def month_index(ts: int) -> int:
    return int(lib().canon_month_index(int(ts)))


def groupby(seg_offsets, dates, categories, cents, seg_lo=0, seg_hi=None):
    """(user, category, month, sum, rows) arrays, ascending key order, non-empty groups only."""
    off = np.ascontiguousarray(seg_offsets, dtype=np.int64)
    seg_hi = len(off) - 1 if seg_hi is None else seg_hi
    d = np.ascontiguousarray(dates, dtype=np.int32)
    c = np.ascontiguousarray(categories, dtype=np.uint8)
    v = np.ascontiguousarray(cents, dtype=np.int32)
    cap = int(off[seg_hi] - off[seg_lo]) + 1
    ou = np.empty(cap, np.int64); oc = np.empty(cap, np.int32); om = np.empty(cap, np.int32)
    os_ = np.empty(cap, np.int64); orow = np.empty(cap, np.int64)
    m = lib().canon_groupby(_ptr(off), seg_lo, seg_hi, _ptr(d), _ptr(c), _ptr(v), cap,
                            _ptr(ou), _ptr(oc), _ptr(om), _ptr(os_), _ptr(orow))
    if m < 0:
        raise RuntimeError(f"canon_groupby failed: {m}")
    return ou[:m].copy(), oc[:m].copy(), om[:m].copy(), os_[:m].copy(), orow[:m].copy()
