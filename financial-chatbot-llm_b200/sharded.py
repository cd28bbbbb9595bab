"""Point-range sharding across the GPUs of one box (BASELINE config 4).

One process per GPU (torch.distributed, NCCL over NVLink/NVSwitch for the plumbing). Rows are
sorted by user and cut into contiguous global row ranges, one per rank; a filter's row range is
clipped by each shard, so a tenant costs nothing on shards it does not touch. Per batch:

  1. every rank runs fcb_search on its shard, writing ids | scores | counts into ONE packed
     int64 buffer (no packing kernel: the three outputs are views of the same allocation);
  2. ONE all_gather_into_tensor of that buffer (batch*k*16 + batch*4 bytes per rank — 10 KB at
     batch 64, k 10: latency-bound on NVSwitch, never bandwidth-bound);
  3. every rank runs the same k-way merge kernel (fcb_merge_topk) -> identical global top-k.

The merge is exact: the global top-k is contained in the union of the per-shard top-k.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Callable, List, Optional, Sequence

import numpy as np

from . import _lib
from .client import B200QdrantClient, Collection
from .qdrant_models import QueryResponse, ScoredPoint


def packed_len(batch: int, k: int) -> int:
    """int64 elements of one rank's packed result buffer."""
    return 2 * batch * k + (batch + 1) // 2


def cuda_merge(device: int, gathered, world: int, batch: int, k: int, out, stream: Optional[int] = None) -> None:
    """fcb_merge_topk over an all-gathered [world, packed_len] int64 CUDA tensor into packed `out`."""
    import torch
    lib = _lib.load()
    P = packed_len(batch, k)
    base = gathered.data_ptr()
    o = out.data_ptr()
    if stream is None:
        stream = torch.cuda.current_stream(device).cuda_stream
    _lib.check(lib.fcb_merge_topk(device, base, base + 8 * batch * k, base + 16 * batch * k, 8 * P, world, batch, k,
                                  o, o + 8 * batch * k, o + 16 * batch * k, C.c_void_p(stream)))


class ShardMerger:
    """Owns the packed local / gathered / merged buffers of one (batch, k) shape."""

    def __init__(self, device, world: int, batch: int, k: int, group: Any = None,
                 merge_impl: Optional[Callable] = None):
        import torch
        self.world, self.batch, self.k, self.group = world, batch, k, group
        self.device = device
        P = packed_len(batch, k)
        self.local = torch.zeros(P, dtype=torch.int64, device=device)
        self.gathered = torch.zeros((world, P), dtype=torch.int64, device=device)
        self.merged = torch.zeros(P, dtype=torch.int64, device=device)
        self._merge = merge_impl

    @staticmethod
    def _views(buf, batch, k):
        import torch
        n = batch * k
        return (buf[:n].view(batch, k), buf[n:2 * n].view(torch.float64).view(batch, k),
                buf[2 * n:].view(torch.int32)[:batch])

    @property
    def local_views(self):
        """(ids int64 [B,k], scores float64 [B,k], counts int32 [B]) — pass as fcb_search outputs."""
        return self._views(self.local, self.batch, self.k)

    @property
    def ids(self):
        return self._views(self.merged, self.batch, self.k)[0]

    @property
    def scores(self):
        return self._views(self.merged, self.batch, self.k)[1]

    @property
    def counts(self):
        return self._views(self.merged, self.batch, self.k)[2]

    def merge(self) -> None:
        """all-gather the local packed results and merge; asynchronous on the current stream."""
        import torch.distributed as dist
        if self.world > 1:
            dist.all_gather_into_tensor(self.gathered.view(-1), self.local, group=self.group)
        else:
            self.gathered[0].copy_(self.local)
        if self._merge is not None:
            self._merge(self.gathered, self.world, self.batch, self.k, self.merged)
        else:
            if not self.gathered.is_cuda:
                raise _lib.FcbError(_lib.FCB_ERR_CUDA, "the shard merge runs on the GPU only (no CPU path)")
            cuda_merge(self.gathered.device.index, self.gathered, self.world, self.batch, self.k, self.merged)


def shard_ranges(n_rows: int, world: int, offsets: Optional[np.ndarray] = None) -> List[tuple]:
    """Contiguous global row ranges, one per rank. With CSR `offsets` the cuts are moved to the
    nearest user boundary (so aggregation cells are disjoint across ranks — no reduce needed)."""
    cuts = [round(n_rows * r / world) for r in range(world + 1)]
    if offsets is not None:
        off = np.asarray(offsets, dtype=np.int64)
        for r in range(1, world):
            j = int(np.searchsorted(off, cuts[r]))
            j = min(max(j, 0), len(off) - 1)
            lo = off[j - 1] if j > 0 else off[0]
            cuts[r] = int(lo if cuts[r] - lo <= off[j] - cuts[r] else off[j])
        for r in range(1, world + 1):
            cuts[r] = max(cuts[r], cuts[r - 1])
    return [(int(cuts[r]), int(cuts[r + 1])) for r in range(world)]


class ShardedSearcher:
    """Collective search over a sharded collection. Call on every rank with identical arguments."""

    def __init__(self, collection: Collection, world: int, device, group: Any = None,
                 merge_impl: Optional[Callable] = None):
        self.col, self.world, self.device, self.group, self._merge_impl = collection, world, device, group, merge_impl
        self._mergers = {}

    def merger(self, batch: int, k: int) -> ShardMerger:
        key = (batch, k)
        if key not in self._mergers:
            self._mergers[key] = ShardMerger(self.device, self.world, batch, k, self.group, self._merge_impl)
        return self._mergers[key]

    def search(self, queries: np.ndarray, filters: Sequence[Any], k: int):
        """queries float64 [B, dim] (host); filters: FcbFilter per query (global rows).
        Returns (ids uint64 [B,k], scores float64 [B,k], counts int32 [B]) — same on every rank."""
        import torch
        q = np.ascontiguousarray(queries, dtype=np.float64).reshape(-1, self.col.engine.dim)
        B = q.shape[0]
        m = self.merger(B, k)
        dq = torch.from_numpy(q).to(self.device)
        ids, scores, counts = m.local_views
        self.col.engine.search_device(dq, list(filters), k, ids, scores, counts)
        m.merge()
        out_ids = m.ids.cpu().numpy().view(np.uint64)
        return out_ids, m.scores.cpu().numpy(), m.counts.cpu().numpy()


class ShardedB200Client(B200QdrantClient):
    """``query_points`` over a sharded collection (collective: every rank makes the same call and
    gets the same answer). Payloads are materialised from the collection's payload function, which
    must be able to serve any global row on every rank."""

    def __init__(self, world: int, device, group: Any = None, merge_impl: Optional[Callable] = None, **kw: Any):
        super().__init__(**kw)
        self._world, self._device, self._group, self._merge_impl = world, device, group, merge_impl
        self._searchers = {}

    def attach_collection(self, collection: Collection) -> None:
        super().attach_collection(collection)
        self._searchers[collection.name] = ShardedSearcher(collection, self._world, self._device, self._group,
                                                           self._merge_impl)

    def query_points(self, collection_name: str, query: Any = None, using=None, prefetch=None, query_filter=None,
                     search_params=None, limit: int = 10, offset: Optional[int] = None, with_payload: Any = True,
                     with_vectors: Any = False, score_threshold: Optional[float] = None, **kwargs: Any) -> QueryResponse:
        col = self.get_collection(collection_name)
        if prefetch is not None or using is not None or with_vectors or query is None:
            raise NotImplementedError("unsupported query_points arguments for the sharded engine")
        q = np.asarray(query, dtype=np.float64)
        skip = int(offset or 0)
        k = int(limit) + skip
        if k < 1:
            return QueryResponse(points=[])
        ids, scores = col.search_union(self._searchers[collection_name].search, q, col.resolve_dnf(query_filter), k)
        pts = []
        for j in range(skip, ids.shape[0]):
            s = float(scores[j])
            if score_threshold is not None and s < score_threshold:
                break
            pid = int(ids[j])
            row = col.row_of_id(pid)                   # KeyError (never a wrong row) if the id map lacks another shard's ids
            pts.append(ScoredPoint(id=col.point_id_fn(row) if col.point_id_fn else pid, version=0, score=s, vector=None,
                                   payload=col.payload_fn(row) if with_payload else None))
        return QueryResponse(points=pts)


# ---- owner-computes routing over NVLink peer memory (fcb_search_routed, csrc/fcb_exchange.cu) ------------------------

FILTER_DTYPE = np.dtype([("row_begin", "<i8"), ("row_end", "<i8"), ("date_lo", "<i8"), ("date_hi", "<i8"),
                         ("category_mask", "<u8")])             # = fcb_filter: batches of filters without a Python loop per call


def filters_array(row_begin, row_end, date_lo=None, date_hi=None, category_mask=None) -> np.ndarray:
    """A batch of fcb_filter as ONE numpy structured array (row ranges as arrays; bounds default to 'no filter')."""
    rb = np.asarray(row_begin, dtype=np.int64)
    out = np.empty(rb.shape[0], dtype=FILTER_DTYPE)
    out["row_begin"], out["row_end"] = rb, np.asarray(row_end, dtype=np.int64)
    out["date_lo"] = _lib.FCB_DATE_MIN if date_lo is None else date_lo
    out["date_hi"] = _lib.FCB_DATE_MAX if date_hi is None else date_hi
    out["category_mask"] = _lib.FCB_ALL_CATEGORIES if category_mask is None else category_mask
    return out


class _DevicePointer:
    """A raw device allocation dressed for torch.as_tensor (zero copy)."""
    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerExchange:
    """The exchange buffers of one (slots, k) shape: this rank's allocation plus every peer's, opened through CUDA IPC
    (the 64-byte handles travel by all_gather_object; after that no collective is involved). `cuts[g] .. cuts[g + 1]` are
    rank g's global rows. One routed search in flight at a time; consume `results()` on the stream of the call before
    the next call."""

    def __init__(self, device: int, world: int, rank: int, slots: int, k: int, cuts: Sequence[int], group: Any = None):
        import torch
        import torch.distributed as dist
        if world > _lib.FCB_MAX_WORLD:
            raise ValueError(f"at most {_lib.FCB_MAX_WORLD} ranks")
        self._lib = _lib.load()
        self.device, self.world, self.rank, self.slots, self.k = device, world, rank, slots, k
        self.bytes = int(self._lib.fcb_exchange_bytes(slots, k))
        own = C.c_void_p()
        handle = (C.c_ubyte * _lib.FCB_PEER_HANDLE_BYTES)()
        _lib.check(self._lib.fcb_peer_alloc(device, self.bytes, C.byref(own), handle))
        handles: List[Any] = [None] * world
        if world > 1:
            dist.all_gather_object(handles, bytes(handle), group=group)
        else:
            handles[0] = bytes(handle)
        self._ptrs: List[int] = []
        for g in range(world):
            if g == rank:
                self._ptrs.append(int(own.value))
                continue
            p = C.c_void_p()
            h = (C.c_ubyte * _lib.FCB_PEER_HANDLE_BYTES).from_buffer_copy(handles[g])
            _lib.check(self._lib.fcb_peer_open(device, h, C.byref(p)))
            self._ptrs.append(int(p.value))
        self._cuts = np.ascontiguousarray(cuts, dtype=np.int64)
        assert self._cuts.shape[0] == world + 1
        bufs = (C.c_void_p * _lib.FCB_MAX_WORLD)(*([C.c_void_p(p) for p in self._ptrs] + [None] * (_lib.FCB_MAX_WORLD - world)))
        self.x = _lib.FcbExchange(world, rank, bufs, self._cuts.ctypes.data)
        self.seq = 0
        self.status = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", device))
        if world > 1:
            dist.barrier(group=group)                  # everybody has opened everybody's buffer before the first push

    def results(self, seq: Optional[int] = None):
        """(ids int64 [slots, k], scores float64 [slots, k], counts int32 [slots]) of exchange `seq` (default: the last
        one) as zero-copy CUDA tensors over this rank's buffer."""
        import torch
        seq = self.seq if seq is None else seq
        base = self._ptrs[self.rank] + (seq & 1) * (self.bytes // 2)
        n = self.slots * self.k
        dev = torch.device("cuda", self.device)
        ids = torch.as_tensor(_DevicePointer(base, (self.slots, self.k), "<i8"), device=dev)
        scores = torch.as_tensor(_DevicePointer(base + 8 * n, (self.slots, self.k), "<f8"), device=dev)
        counts = torch.as_tensor(_DevicePointer(base + 16 * n, (self.slots,), "<i4"), device=dev)
        return ids, scores, counts

    def close(self) -> None:
        for g, p in enumerate(self._ptrs):
            self._lib.fcb_peer_close(self.device, C.c_void_p(p), 1 if g == self.rank else 0)
        self._ptrs = []


class RoutedSearcher:
    """Collective search where every tenant lives on ONE shard (cuts on user boundaries): the owner computes the final
    top-k of its queries and writes it into every rank's exchange buffer; no all-gather, no merge. Call on every rank
    with the same queries (resident on every rank's GPU), the same filters and the same k."""

    def __init__(self, collection: Collection, exchange: PeerExchange):
        self.col, self.xch = collection, exchange
        self._lib = _lib.load()

    def search_device(self, d_queries, filters: np.ndarray, k: int, stream: Optional[int] = None) -> int:
        """Asynchronous. `filters`: structured array (FILTER_DTYPE), one per exchange slot. Returns the sequence number;
        `exchange.results(seq)` holds the answers once the stream has run the call."""
        import torch
        xch = self.xch
        if filters.dtype != FILTER_DTYPE or filters.shape[0] != xch.slots or k != xch.k:
            raise ValueError("filters must be a FILTER_DTYPE array with one entry per exchange slot, k the exchange's k")
        if stream is None:
            stream = torch.cuda.current_stream(xch.device).cuda_stream
        xch.seq += 1
        eng = self.col.engine
        _lib.check(self._lib.fcb_search_routed(eng._h, C.byref(xch.x), d_queries.data_ptr(), filters.ctypes.data, xch.slots, k,
                                               xch.seq, xch.status.data_ptr(), C.c_void_p(stream)), eng._h)
        return xch.seq

    def search(self, queries: np.ndarray, filters: np.ndarray, k: int):
        """Host arrays in, host arrays out (ids uint64, scores, counts); synchronous."""
        import torch
        dev = torch.device("cuda", self.xch.device)
        dq = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float64)).to(dev)
        seq = self.search_device(dq, filters, k)
        ids, scores, counts = self.xch.results(seq)
        out = ids.cpu().numpy().view(np.uint64), scores.cpu().numpy(), counts.cpu().numpy()
        if int(self.xch.status.item()) != 0:
            raise _lib.FcbError(_lib.FCB_ERR_STATE, "routed search: a peer never delivered its results (exchange timed out)")
        return out
