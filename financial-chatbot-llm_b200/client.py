"""Drop-in stand-in for ``qdrant_client.QdrantClient`` on the hot path.

The reference builds its backend once per process at tools/qdrant_tool.py:24-27
(``QdrantHttpClient(url=QDRANT_URL, api_key=QDRANT_API_KEY)``) and calls
``client.query_points(collection_name=, query=, limit=, search_params=,
query_filter=).points`` at :147-153, reading ``hit.payload`` at :159-164.
``B200QdrantClient`` accepts the same constructor keywords (ignored: there is no
server) and the same ``query_points`` call, and returns the same shape — but the
search runs in libfcb200.so on a B200 over a resident, CSR-by-user corpus.

Filter subset (anything else raises NotImplementedError — never a silent wrong answer):
``Filter(must=[...])`` of ``FieldCondition`` on
  metadata.user_id  MatchValue(str)             -> the user's CSR segment
  metadata.date     Range(gte/gt/lte/lt)        -> inclusive int64 bounds fused into the scan
  metadata.category MatchValue / MatchAny       -> 64-bit category mask fused into the scan
Filters are read by attribute, so real qdrant-client model instances work too.
"""
from __future__ import annotations

import math
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from . import _lib
from .engine import Engine, make_filter
from .qdrant_models import QueryResponse, ScoredPoint

USER_KEY, DATE_KEY, CATEGORY_KEY = "metadata.user_id", "metadata.date", "metadata.category"


class ParsedFilter:
    __slots__ = ("user_id", "has_user", "date_lo", "date_hi", "category_values")

    def __init__(self):
        self.user_id: Optional[str] = None
        self.has_user = False
        self.date_lo: Optional[int] = None
        self.date_hi: Optional[int] = None
        self.category_values: Optional[set] = None     # None = no category filter


def _range_to_bounds(rng: Any):
    lo = hi = None
    def upd_lo(v):
        nonlocal lo
        lo = v if lo is None else max(lo, v)
    def upd_hi(v):
        nonlocal hi
        hi = v if hi is None else min(hi, v)
    if getattr(rng, "gte", None) is not None:
        upd_lo(math.ceil(rng.gte))
    if getattr(rng, "gt", None) is not None:
        upd_lo(math.floor(rng.gt) + 1)
    if getattr(rng, "lte", None) is not None:
        upd_hi(math.floor(rng.lte))
    if getattr(rng, "lt", None) is not None:
        upd_hi(math.ceil(rng.lt) - 1)
    return lo, hi


def parse_filter(flt: Any) -> ParsedFilter:
    """Resolve the qdrant filter model into the column predicate the kernels fuse."""
    out = ParsedFilter()
    if flt is None:
        return out
    for name in ("should", "must_not", "min_should"):
        if getattr(flt, name, None):
            raise NotImplementedError(f"Filter.{name} is not supported by the B200 engine")
    for cond in (getattr(flt, "must", None) or []):
        key = getattr(cond, "key", None)
        match, rng = getattr(cond, "match", None), getattr(cond, "range", None)
        if key is None or (match is None) == (rng is None):
            raise NotImplementedError(f"unsupported condition {cond!r}")
        if key == USER_KEY and match is not None and hasattr(match, "value"):
            v = match.value
            if out.has_user and out.user_id != v:
                out.user_id = None            # two different users ANDed: nothing matches
            else:
                out.user_id = v if isinstance(v, str) else None
            out.has_user = True
        elif key == DATE_KEY and rng is not None:
            lo, hi = _range_to_bounds(rng)
            if lo is not None:
                out.date_lo = lo if out.date_lo is None else max(out.date_lo, lo)
            if hi is not None:
                out.date_hi = hi if out.date_hi is None else min(out.date_hi, hi)
        elif key == CATEGORY_KEY and match is not None:
            vals = {match.value} if hasattr(match, "value") else set(getattr(match, "any"))
            out.category_values = vals if out.category_values is None else (out.category_values & vals)
        else:
            raise NotImplementedError(f"unsupported filter condition on key {key!r}")
    return out


class Collection:
    """Host-side description of one resident collection (one shard on one GPU).

    users        sorted list of user_id strings; user u owns global rows
                 [offsets[u], offsets[u+1]) (CSR, rows sorted by user)
    payload_fn   global row -> payload dict ({"page_content": str, "metadata": {...}}), only
                 called for the k winners
    category_codes  payload category value -> code 0..63 used in the device column
    """

    def __init__(self, name: str, engine: Engine, users: Sequence[str], offsets: np.ndarray,
                 payload_fn: Callable[[int], Dict[str, Any]], *, n_total_rows: Optional[int] = None,
                 category_codes: Optional[Dict[Any, int]] = None, ids_host: Optional[np.ndarray] = None):
        self.name, self.engine = name, engine
        self.users = list(users)
        self.user_index = {u: i for i, u in enumerate(self.users)}
        self.offsets = np.asarray(offsets, dtype=np.int64)
        self.payload_fn = payload_fn
        self.n_total_rows = int(self.offsets[-1]) if n_total_rows is None else n_total_rows
        self.category_codes = category_codes or {}
        self.point_id_fn: Optional[Callable[[int], Any]] = None   # global row -> original point id (non-integer ids)
        if ids_host is not None:
            ids_host = np.asarray(ids_host, dtype=np.uint64)
            self._id_order = np.argsort(ids_host, kind="stable")
            self._ids_sorted = ids_host[self._id_order]
        else:
            self._id_order = self._ids_sorted = None

    def row_of_id(self, pid: int) -> int:
        """Global row of a returned point id."""
        if self._ids_sorted is None:
            return int(pid)
        j = int(np.searchsorted(self._ids_sorted, np.uint64(pid)))
        return int(self._id_order[j]) + self.engine.row_base

    def resolve(self, pf: ParsedFilter):
        """ParsedFilter -> FcbFilter for this collection."""
        if pf.has_user:
            u = self.user_index.get(pf.user_id) if pf.user_id is not None else None
            rb, re_ = (0, 0) if u is None else (int(self.offsets[u]), int(self.offsets[u + 1]))
        else:
            rb, re_ = 0, self.n_total_rows
        mask = None
        if pf.category_values is not None:
            mask = 0
            for v in pf.category_values:
                c = self.category_codes.get(v)
                if c is not None and 0 <= c < 64:
                    mask |= 1 << c
        return make_filter(rb, re_, pf.date_lo, pf.date_hi, mask)


class B200QdrantClient:
    """``QdrantClient`` look-alike whose ``query_points`` runs on the B200 engine."""

    def __init__(self, location: Optional[str] = None, url: Optional[str] = None,
                 api_key: Optional[str] = None, **_ignored: Any):
        # url / api_key name a remote server in the reference (config.py:45-46); nothing is remote here
        self._collections: Dict[str, Collection] = {}

    def attach_collection(self, collection: Collection) -> None:
        self._collections[collection.name] = collection

    def get_collection(self, collection_name: str) -> Collection:
        try:
            return self._collections[collection_name]
        except KeyError:
            raise ValueError(f"Collection {collection_name} not found") from None

    def query_points(self, collection_name: str, query: Any = None, using: Optional[str] = None,
                     prefetch: Any = None, query_filter: Any = None, search_params: Any = None,
                     limit: int = 10, offset: Optional[int] = None, with_payload: Any = True,
                     with_vectors: Any = False, score_threshold: Optional[float] = None,
                     **kwargs: Any) -> QueryResponse:
        col = self.get_collection(collection_name)
        if prefetch is not None or using is not None:
            raise NotImplementedError("prefetch / named vectors are not supported")
        if with_vectors:
            raise NotImplementedError("with_vectors=True is not supported")
        if query is None:
            raise NotImplementedError("query_points without a query vector (scroll) is not supported")
        # search_params (hnsw_ef, exact) steer an approximate index; this engine is always exact
        q = np.asarray(query, dtype=np.float64)
        if q.ndim != 1 or q.shape[0] != col.engine.dim:
            raise ValueError(f"query must be a vector of length {col.engine.dim}")
        skip = int(offset or 0)
        k = int(limit) + skip
        if k < 1:
            return QueryResponse(points=[])
        filt = col.resolve(parse_filter(query_filter))
        ids, scores, counts = col.engine.search_host(q[None, :], [filt], k)
        n = int(counts[0])
        points: List[ScoredPoint] = []
        for j in range(skip, n):
            s = float(scores[0, j])
            if score_threshold is not None and s < score_threshold:
                break
            pid = int(ids[0, j])
            row = col.row_of_id(pid)
            payload = col.payload_fn(row) if with_payload else None
            points.append(ScoredPoint(id=col.point_id_fn(row) if col.point_id_fn else pid, version=0, score=s,
                                      payload=payload, vector=None))
        return QueryResponse(points=points)
