"""Thin object wrapper over the C-ABI (one engine per process per GPU).

PyTorch is used only to hold device memory: every tensor handed to the library is
passed as ``tensor.data_ptr()`` and kept referenced here so it outlives the shard.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import FcbConfig, FcbFilter, FcbShard, check


def make_filter(row_begin: int, row_end: int, date_lo: Optional[int] = None,
                date_hi: Optional[int] = None, category_mask: Optional[int] = None) -> FcbFilter:
    return FcbFilter(int(row_begin), int(row_end),
                     _lib.FCB_DATE_MIN if date_lo is None else max(int(date_lo), _lib.FCB_DATE_MIN),
                     _lib.FCB_DATE_MAX if date_hi is None else min(int(date_hi), _lib.FCB_DATE_MAX),
                     _lib.FCB_ALL_CATEGORIES if category_mask is None else int(category_mask))


class Engine:
    def __init__(self, device: int = 0, max_batch: int = 1, max_k: int = _lib.FCB_MAX_K):
        """``max_k`` defaults to FCB_MAX_K = 10000: the reference's default call (num_transactions=None) asks for
        limit=10000 (tools/qdrant_tool.py:145); an engine opened with a smaller max_k refuses that call, and the
        tool's blanket ``except`` would turn the refusal into ``[]``. Costs 8 bytes of scratch per shard row."""
        self._lib = _lib.load()
        self._h = C.c_void_p()
        cfg = FcbConfig(device, max_batch, max_k, 0)
        check(self._lib.fcb_open(C.byref(cfg), C.byref(self._h)))
        self.device, self.max_batch, self.max_k = device, max_batch, max_k
        self._keep = None
        self.dim = 0
        self.n_rows = 0
        self.row_base = 0

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.fcb_close(self._h)
            self._h = C.c_void_p()
            self._keep = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- shard
    def load_shard(self, vectors, *, ids=None, id_rank=None, dates=None, categories=None, cents=None,
                   seg_offsets=None, row_base: int = 0) -> None:
        """``vectors``: CUDA tensor [n, dim], float32 or bfloat16, contiguous. Optional CUDA
        column tensors: ids uint64/int64 [n], id_rank int32/uint32 [n], dates int32 [n],
        categories uint8 [n], cents int32 [n], seg_offsets int64 [n_segs+1]."""
        import torch
        if not vectors.is_cuda or not vectors.is_contiguous() or vectors.dim() != 2:
            raise ValueError("vectors must be a contiguous 2-D CUDA tensor")
        if vectors.dtype == torch.float32:
            dt = _lib.FCB_DTYPE_F32
        elif vectors.dtype == torch.bfloat16:
            dt = _lib.FCB_DTYPE_BF16
        else:
            raise ValueError("vectors must be float32 or bfloat16")
        n, dim = vectors.shape

        def col(t, dtypes, name, length=n):
            if t is None:
                return None
            if not t.is_cuda or not t.is_contiguous() or t.dtype not in dtypes or t.numel() != length:
                raise ValueError(f"{name}: need contiguous CUDA tensor of {dtypes} with {length} elements")
            return t.data_ptr()
        n_segs = 0 if seg_offsets is None else seg_offsets.numel() - 1
        sh = FcbShard(vectors.data_ptr(),
                      col(ids, (torch.int64, torch.uint64), "ids"),
                      col(id_rank, (torch.int32, torch.uint32), "id_rank"),
                      col(dates, (torch.int32,), "dates"),
                      col(categories, (torch.uint8,), "categories"),
                      col(cents, (torch.int32,), "cents"),
                      col(seg_offsets, (torch.int64,), "seg_offsets", n_segs + 1) if seg_offsets is not None else None,
                      n, n_segs, row_base, dim, dt)
        check(self._lib.fcb_load_shard(self._h, C.byref(sh)), self._h)
        self._keep = (vectors, ids, id_rank, dates, categories, cents, seg_offsets)
        self.dim, self.n_rows, self.row_base = dim, n, row_base

    # ---- search
    def search_host(self, queries: np.ndarray, filters: Sequence[FcbFilter], k: int):
        """queries float64 [B, dim] (host). Returns (ids uint64 [B,k], scores float64 [B,k],
        counts int32 [B]) as numpy arrays. Synchronous (fcb_search_host)."""
        q = np.ascontiguousarray(queries, dtype=np.float64).reshape(-1, self.dim)
        B = q.shape[0]
        if len(filters) != B:
            raise ValueError("one filter per query")
        farr = (FcbFilter * B)(*filters)
        ids = np.empty((B, k), dtype=np.uint64)
        scores = np.empty((B, k), dtype=np.float64)
        counts = np.empty(B, dtype=np.int32)
        ptr = lambda a: a.__array_interface__["data"][0]         # several times cheaper than ndarray.ctypes.data
        check(self._lib.fcb_search_host(self._h, ptr(q), farr, B, k, ptr(ids), ptr(scores), ptr(counts)), self._h)
        return ids, scores, counts

    def search_device(self, d_queries, filters: Sequence[FcbFilter], k: int, out_ids, out_scores,
                      out_counts, stream: Optional[int] = None) -> None:
        """Asynchronous fcb_search: CUDA tensors in (float64 [B,dim]) and out (int64/uint64 [B,k],
        float64 [B,k], int32 [B]); ``stream`` is a raw cudaStream_t (default: torch's current)."""
        import torch
        B = d_queries.shape[0]
        farr = (FcbFilter * B)(*filters)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._lib.fcb_search(self._h, d_queries.data_ptr(), farr, B, k, out_ids.data_ptr(),
                                   out_scores.data_ptr(), out_counts.data_ptr(), C.c_void_p(stream)), self._h)

    # ---- aggregation
    def aggregate(self, seg_begin: int, seg_end: int, month_lo: int, n_months: int, n_categories: int,
                  sums=None, present=None, dropped=None, stream: Optional[int] = None):
        """fcb_aggregate over the loaded shard. Returns CUDA tensors (sums int64 [S, C, M],
        present uint8 [S, C, M], dropped int64 [1]); asynchronous on torch's current stream."""
        import torch
        dev = torch.device("cuda", self.device)
        S = seg_end - seg_begin
        if sums is None:
            sums = torch.empty((S, n_categories, n_months), dtype=torch.int64, device=dev)
        if present is None:
            present = torch.empty((S, n_categories, n_months), dtype=torch.uint8, device=dev)
        if dropped is None:
            dropped = torch.empty(1, dtype=torch.int64, device=dev)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._lib.fcb_aggregate(self._h, seg_begin, seg_end, month_lo, n_months, n_categories,
                                      sums.data_ptr(), present.data_ptr(), dropped.data_ptr(),
                                      C.c_void_p(stream)), self._h)
        return sums, present, dropped

    def aggregate_hits(self, hit_ids, hit_counts, month_lo: int, n_months: int, n_categories: int,
                       stream: Optional[int] = None):
        """fcb_aggregate_hits: (category, month) sums over the points a search returned. ``hit_ids`` int64/uint64
        CUDA [B, k] and ``hit_counts`` int32 CUDA [B] as written by ``search_device``. Returns CUDA tensors
        (sums int64 [B, C, M], present uint8 [B, C, M], stats int64 [2]); asynchronous on torch's current stream."""
        import torch
        dev = torch.device("cuda", self.device)
        B, k = hit_ids.shape
        sums = torch.empty((B, n_categories, n_months), dtype=torch.int64, device=dev)
        present = torch.empty((B, n_categories, n_months), dtype=torch.uint8, device=dev)
        stats = torch.empty(2, dtype=torch.int64, device=dev)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        check(self._lib.fcb_aggregate_hits(self._h, hit_ids.data_ptr(), hit_counts.data_ptr(), B, k, month_lo, n_months,
                                           n_categories, sums.data_ptr(), present.data_ptr(), stats.data_ptr(),
                                           C.c_void_p(stream)), self._h)
        return sums, present, stats

    @property
    def scan_config(self):
        """(rows per warp step, unroll) of the fused scan measured fastest on the loaded shard; (0, 0) = table default."""
        r, u = C.c_int32(0), C.c_int32(0)
        check(self._lib.fcb_scan_config(self._h, C.byref(r), C.byref(u)), self._h)
        return int(r.value), int(u.value)

    @property
    def last_search_ms(self) -> float:
        return float(self._lib.fcb_last_search_ms(self._h))

    @property
    def dates_sorted(self) -> bool:
        """True if the loaded shard's dates are non-decreasing within every user (atomic-free aggregation)."""
        return int(self._lib.fcb_shard_dates_sorted(self._h)) == 1

    def last_dense_redo(self, batch: int) -> int:
        """Queries of the last batched tensor-core call answered by the exact re-run (-1: path inactive)."""
        return int(self._lib.fcb_last_dense_redo(self._h, batch))

    def last_dense_widened(self, batch: int) -> int:
        """Queries of the last batched tensor-core call answered by re-scoring more than the 32 best listed rows."""
        return int(self._lib.fcb_last_dense_widened(self._h, batch))

    @property
    def exact_reruns(self) -> int:
        """Queries since load_shard that a fast path could not certify and the exact float64 kernel answered."""
        return int(self._lib.fcb_exact_reruns(self._h))

    @property
    def last_search_launches(self) -> int:
        return int(self._lib.fcb_last_search_launches(self._h))


def build_group_keys(device: int, dates, categories, cents, stream: Optional[int] = None):
    """fcb_group_keys_build: the 16-bit (category, month) key of every row of caller-owned CUDA columns, resolved once so
    that ``aggregate_columns(..., keys=, key_base=)`` reads 6 bytes per row instead of 9. Returns (keys uint16-as-int16
    CUDA tensor [n_rows], key_base). Waits for the stream."""
    import ctypes
    import torch
    lib = _lib.load()
    n = int(cents.numel())
    keys = torch.empty(((n + 7) // 8) * 8, dtype=torch.int16, device=torch.device("cuda", device))
    base = ctypes.c_int32(0)
    if stream is None:
        stream = torch.cuda.current_stream(device).cuda_stream
    check(lib.fcb_group_keys_build(device, dates.data_ptr(), categories.data_ptr(), cents.data_ptr(), n, keys.data_ptr(),
                                   ctypes.byref(base), stream))
    return keys[:n], int(base.value)


def aggregate_columns(device: int, dates, categories, cents, seg_offsets, seg_begin: int, seg_end: int, month_lo: int,
                      n_months: int, n_categories: int, sums=None, present=None, dropped=None, *, row_base: int = 0,
                      dates_sorted: bool = False, stream: Optional[int] = None, keys=None, key_base: int = 0):
    """fcb_aggregate_columns(_keyed): the (user, category, month) integer-cents group-by over caller-owned CUDA columns,
    without an engine (no vectors, nothing allocated inside the library). ``keys`` / ``key_base`` as returned by
    ``build_group_keys`` (``dates`` / ``categories`` may then be None if no window ever leaves the keys' range). Returns
    (sums int64 [S, C, M], present uint8 [S, C, M], dropped int64 [1]); asynchronous on torch's current stream."""
    import torch
    lib = _lib.load()
    dev = torch.device("cuda", device)
    S = seg_end - seg_begin
    if sums is None:
        sums = torch.empty((S, n_categories, n_months), dtype=torch.int64, device=dev)
    if present is None:
        present = torch.empty((S, n_categories, n_months), dtype=torch.uint8, device=dev)
    if dropped is None:
        dropped = torch.empty(1, dtype=torch.int64, device=dev)
    if stream is None:
        stream = torch.cuda.current_stream(device).cuda_stream
    ptr = lambda x: None if x is None else x.data_ptr()
    if keys is None:
        check(lib.fcb_aggregate_columns(device, dates.data_ptr(), categories.data_ptr(), cents.data_ptr(), seg_offsets.data_ptr(),
                                        cents.numel(), seg_offsets.numel() - 1, row_base, 1 if dates_sorted else 0, seg_begin,
                                        seg_end, month_lo, n_months, n_categories, sums.data_ptr(), present.data_ptr(),
                                        dropped.data_ptr(), C.c_void_p(stream)))
    else:
        check(lib.fcb_aggregate_columns_keyed(device, keys.data_ptr(), key_base, ptr(dates), ptr(categories), cents.data_ptr(),
                                              seg_offsets.data_ptr(), cents.numel(), seg_offsets.numel() - 1, row_base,
                                              1 if dates_sorted else 0, seg_begin, seg_end, month_lo, n_months, n_categories,
                                              sums.data_ptr(), present.data_ptr(), dropped.data_ptr(), C.c_void_p(stream)))
    return sums, present, dropped


def synth_vectors(out, row_lo: int, seed: int, n_clusters: int) -> None:
    """Fill CUDA tensor ``out`` [n, dim] (float32/bfloat16) with rows row_lo.. of the synthetic corpus."""
    import torch
    lib = _lib.load()
    dt = _lib.FCB_DTYPE_F32 if out.dtype == torch.float32 else _lib.FCB_DTYPE_BF16
    n, dim = out.shape
    st = torch.cuda.current_stream(out.device).cuda_stream
    check(lib.fcb_synth_vectors(out.device.index, out.data_ptr(), dt, row_lo, n, dim, seed, n_clusters,
                                C.c_void_p(st)))


def synth_columns(row_lo: int, n_rows: int, seed: int, n_categories: int, date_span_seconds: int,
                  dates=None, categories=None, cents=None) -> None:
    import torch
    lib = _lib.load()
    t = next(x for x in (dates, categories, cents) if x is not None)
    st = torch.cuda.current_stream(t.device).cuda_stream
    p = lambda x: None if x is None else x.data_ptr()
    check(lib.fcb_synth_columns(t.device.index, p(dates), p(categories), p(cents), row_lo, n_rows, seed,
                                n_categories, date_span_seconds, C.c_void_p(st)))


def search_noalloc(device: int, shard: FcbShard, d_query, filt: FcbFilter, k: int, out_ids, out_scores,
                   out_count, stream: Optional[int] = None) -> None:
    """fcb_search_noalloc: one query, no engine and no allocation inside the library (for callers
    that cannot allocate GPU memory — see include/fcb200.h). ``shard`` is a host FcbShard whose
    pointers are CUDA tensors' data_ptr(); query/outputs are CUDA tensors the caller owns.
    Asynchronous on ``stream`` (default: torch's current stream). Calls sharing a GPU must not overlap."""
    import torch
    lib = _lib.load()
    if stream is None:
        stream = torch.cuda.current_stream(device).cuda_stream
    err = C.create_string_buffer(256)
    rc = lib.fcb_search_noalloc(device, C.byref(shard), d_query.data_ptr(), C.byref(filt), k,
                                out_ids.data_ptr(), out_scores.data_ptr(), out_count.data_ptr(),
                                C.c_void_p(stream), err, 256)
    if rc != _lib.FCB_OK:
        raise _lib.FcbError(rc, err.value.decode() or "fcb_search_noalloc failed")


def synth_dates_sorted(dates, seg_offsets, row_lo: int, date_span_seconds: int) -> None:
    """Fill CUDA int32 tensor ``dates`` (rows row_lo..) with time-ordered dates per user segment
    (``seg_offsets``: CUDA int64 CSR offsets in global rows)."""
    import torch
    lib = _lib.load()
    st = torch.cuda.current_stream(dates.device).cuda_stream
    check(lib.fcb_synth_dates_sorted(dates.device.index, dates.data_ptr(), seg_offsets.data_ptr(),
                                     seg_offsets.numel() - 1, row_lo, dates.numel(), date_span_seconds, C.c_void_p(st)))
