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
