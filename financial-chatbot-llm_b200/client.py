"""Drop-in stand-in for ``qdrant_client.QdrantClient`` on the hot path.

The reference builds its backend once per process at tools/qdrant_tool.py:24-27
(``QdrantHttpClient(url=QDRANT_URL, api_key=QDRANT_API_KEY)``) and calls
``client.query_points(collection_name=, query=, limit=, search_params=,
query_filter=).points`` at :147-153, reading ``hit.payload`` at :159-164.
``B200QdrantClient`` accepts the same constructor keywords (ignored: there is no
server) and the same ``query_points`` call, and returns the same shape — but the
search runs in libfcb200.so on a B200 over a resident, CSR-by-user corpus.

Filter subset (anything else raises NotImplementedError — never a silent wrong answer):
``Filter(must=[...], should=[...], must_not=[...])`` of ``FieldCondition`` on
  metadata.user_id  MatchValue(str)             -> the user's CSR segment
  metadata.date     Range(gte/gt/lte/lt)        -> inclusive int64 bounds fused into the scan
  metadata.category MatchValue / MatchAny       -> 64-bit category mask fused into the scan
The device predicate is one conjunction (row range AND date range AND category mask). ``should`` (OR)
and ``must_not`` (NOT) are rewritten on the host into a disjunction of such conjunctions — NOT of a
date range is "before it OR after it", NOT of a user is "the rows before OR after the segment", NOT of
a category set clears mask bits — every conjunct rides in the same grouped ``fcb_search`` call with a
copy of the query, and the per-conjunct top-k lists are merged (duplicates dropped by point id; the
scores of one row are bit-identical in every list). ``min_should`` and nested filters raise.
Filters are read by attribute, so real qdrant-client model instances work too.
"""
from __future__ import annotations

import bisect
import math
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from . import _lib
from .engine import Engine, make_filter
from .qdrant_models import QueryResponse, ScoredPoint

USER_KEY, DATE_KEY, CATEGORY_KEY = "metadata.user_id", "metadata.date", "metadata.category"
MAX_CONJUNCTS = 64                      # a should / must_not rewrite larger than this is refused


class ParsedFilter:
    """One conjunction, still symbolic (no collection needed): user == / < / >, date bounds, category sets."""
    __slots__ = ("user_id", "has_user", "user_lt", "user_gt", "date_lo", "date_hi", "category_values",
                 "category_excluded")

    def __init__(self):
        self.user_id: Optional[str] = None
        self.has_user = False
        self.user_lt: Optional[str] = None             # only users sorting strictly before this one
        self.user_gt: Optional[str] = None             # only users sorting strictly after this one
        self.date_lo: Optional[int] = None
        self.date_hi: Optional[int] = None
        self.category_values: Optional[set] = None     # None = no category filter
        self.category_excluded: set = set()

    def copy(self) -> "ParsedFilter":
        c = ParsedFilter()
        for f in self.__slots__:
            v = getattr(self, f)
            setattr(c, f, set(v) if isinstance(v, set) else v)
        return c

    def impossible(self) -> bool:
        return (self.has_user and self.user_id is None) or \
               (self.date_lo is not None and self.date_hi is not None and self.date_lo > self.date_hi) or \
               (self.category_values is not None and not (self.category_values - self.category_excluded))


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


def _classify(cond: Any):
    """-> ("user", str | None) | ("date", (lo, hi)) | ("category", set)"""
    key = getattr(cond, "key", None)
    match, rng = getattr(cond, "match", None), getattr(cond, "range", None)
    if key is None or (match is None) == (rng is None):
        raise NotImplementedError(f"unsupported condition {cond!r}")
    if key == USER_KEY and match is not None and hasattr(match, "value"):
        return "user", (match.value if isinstance(match.value, str) else None)
    if key == DATE_KEY and rng is not None:
        return "date", _range_to_bounds(rng)
    if key == CATEGORY_KEY and match is not None:
        return "category", ({match.value} if hasattr(match, "value") else set(getattr(match, "any")))
    raise NotImplementedError(f"unsupported filter condition on key {key!r}")


def _apply(pf: ParsedFilter, kind: str, arg: Any, negate: bool) -> List[ParsedFilter]:
    """pf AND cond (or AND NOT cond) -> the conjuncts it becomes (0, 1 or 2)."""
    out = pf.copy()
    if kind == "user":
        if not negate:
            if arg is None or (out.has_user and out.user_id != arg):
                out.user_id = None                     # non-string value / two different users ANDed: nothing matches
            else:
                out.user_id = arg
            out.has_user = True
            return [out]
        if arg is None:                                # NOT (user == <non-string>): every row passes
            return [out]
        before, after = out, pf.copy()
        before.user_lt = arg if before.user_lt is None else min(before.user_lt, arg)
        after.user_gt = arg if after.user_gt is None else max(after.user_gt, arg)
        return [before, after]
    if kind == "date":
        lo, hi = arg
        if not negate:
            if lo is not None:
                out.date_lo = lo if out.date_lo is None else max(out.date_lo, lo)
            if hi is not None:
                out.date_hi = hi if out.date_hi is None else min(out.date_hi, hi)
            return [out]
        alts = []
        if lo is not None:                             # strictly before the range
            a = pf.copy()
            a.date_hi = lo - 1 if a.date_hi is None else min(a.date_hi, lo - 1)
            alts.append(a)
        if hi is not None:                             # strictly after it
            a = pf.copy()
            a.date_lo = hi + 1 if a.date_lo is None else max(a.date_lo, hi + 1)
            alts.append(a)
        return alts                                    # an unbounded range matches every dated row: NOT of it, none
    if not negate:
        out.category_values = set(arg) if out.category_values is None else (out.category_values & arg)
    else:
        out.category_excluded |= arg
    return [out]


def parse_filter_dnf(flt: Any) -> List[ParsedFilter]:
    """The qdrant filter as a list of conjunctions whose UNION is the set of passing points
    (qdrant semantics: every ``must``, at least one ``should`` when any is given, no ``must_not``)."""
    if flt is None:
        return [ParsedFilter()]
    if getattr(flt, "min_should", None):
        raise NotImplementedError("Filter.min_should is not supported by the B200 engine")
    conj = [ParsedFilter()]
    for cond in (getattr(flt, "must", None) or []):
        kind, arg = _classify(cond)
        conj = [c for pf in conj for c in _apply(pf, kind, arg, False)]
    for cond in (getattr(flt, "must_not", None) or []):
        kind, arg = _classify(cond)
        conj = [c for pf in conj for c in _apply(pf, kind, arg, True) if not c.impossible()]
        if len(conj) > MAX_CONJUNCTS:
            raise NotImplementedError("must_not expands into too many conjunctions")
    should = [_classify(c) for c in (getattr(flt, "should", None) or [])]
    if should:
        conj = [c for pf in conj for kind, arg in should for c in _apply(pf, kind, arg, False)]
        if len(conj) > MAX_CONJUNCTS:
            raise NotImplementedError("should x must_not expands into too many conjunctions")
    return [c for c in conj if not c.impossible()]


def parse_filter(flt: Any) -> ParsedFilter:
    """The filter as ONE conjunction (what the reference's own filter always is, tools/qdrant_tool.py:98-131);
    raises NotImplementedError when it needs more than one — use ``parse_filter_dnf`` for those."""
    conj = parse_filter_dnf(flt)
    if len(conj) > 1:
        raise NotImplementedError("this filter needs several conjunctions; use parse_filter_dnf")
    if not conj:                                       # nothing can match: an empty user segment says so
        pf = ParsedFilter()
        pf.has_user = True
        return pf
    return conj[0]


def merge_conjunct_lists(ids: np.ndarray, scores: np.ndarray, counts: np.ndarray, k: int):
    """Top-k of the UNION of per-conjunct top-k lists (rows of ids / scores, ``counts`` valid entries each):
    duplicates dropped by id, order (score desc, id asc) as everywhere else. -> (ids[n], scores[n]), n <= k."""
    all_ids = np.concatenate([ids[b, :int(counts[b])] for b in range(ids.shape[0])])
    all_sc = np.concatenate([scores[b, :int(counts[b])] for b in range(ids.shape[0])])
    all_ids, first = np.unique(all_ids, return_index=True)
    all_sc = all_sc[first]
    order = np.lexsort((all_ids, -all_sc))[:k]
    return all_ids[order], all_sc[order]


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
                 category_codes: Optional[Dict[Any, int]] = None, ids_host: Optional[np.ndarray] = None,
                 ids_row_base: Optional[int] = None):
        self.name, self.engine = name, engine
        self.users = list(users)
        self.user_index = {u: i for i, u in enumerate(self.users)}
        self.offsets = np.asarray(offsets, dtype=np.int64)
        self.payload_fn = payload_fn
        self.n_total_rows = int(self.offsets[-1]) if n_total_rows is None else n_total_rows
        self.category_codes = category_codes or {}
        self.point_id_fn: Optional[Callable[[int], Any]] = None   # global row -> original point id (non-integer ids)
        self.has_missing_category = False              # True when some rows carry no category (code >= 64 on the device)
        self.has_missing_date = False                  # True when some rows carry no date (stored as INT32_MIN, corpus_io.NO_DATE)
        # ids_host[i] = point id of GLOBAL row ids_row_base + i (default: this engine's shard). A sharded
        # collection needs the ids of EVERY shard here: the merged answer names points of other ranks too.
        self._ids_row_base = int(getattr(engine, "row_base", 0) if ids_row_base is None else ids_row_base)
        if ids_host is not None:
            ids_host = np.asarray(ids_host, dtype=np.uint64)
            self._id_order = np.argsort(ids_host, kind="stable")
            self._ids_sorted = ids_host[self._id_order]
        else:
            self._id_order = self._ids_sorted = None

    def row_of_id(self, pid: int) -> int:
        """Global row of a returned point id. Raises KeyError for an id this collection's id map does not
        hold (never a neighbouring row: that would hand out another transaction's payload)."""
        if self._ids_sorted is None:
            row = int(pid)
            if not 0 <= row < self.n_total_rows:
                raise KeyError(f"point id {pid} is not a row of collection {self.name!r}")
            return row
        j = int(np.searchsorted(self._ids_sorted, np.uint64(pid)))
        if j >= self._ids_sorted.shape[0] or int(self._ids_sorted[j]) != int(pid):
            raise KeyError(f"point id {pid} is not in the id map of collection {self.name!r} "
                           f"(a sharded collection needs the ids of all shards: ids_host=<all ids>, ids_row_base=0)")
        return int(self._id_order[j]) + self._ids_row_base

    def _mask_of(self, values) -> int:
        mask = 0
        for v in values:
            c = self.category_codes.get(v)
            if c is not None and 0 <= c < 64:
                mask |= 1 << c
        return mask

    def resolve(self, pf: ParsedFilter):
        """ParsedFilter -> FcbFilter for this collection."""
        if pf.has_user:
            u = self.user_index.get(pf.user_id) if pf.user_id is not None else None
            rb, re_ = (0, 0) if u is None else (int(self.offsets[u]), int(self.offsets[u + 1]))
        else:
            rb, re_ = 0, self.n_total_rows
        if pf.user_gt is not None:                     # users are sorted: "after user X" is a row suffix
            rb = max(rb, int(self.offsets[bisect.bisect_right(self.users, pf.user_gt)]))
        if pf.user_lt is not None:
            re_ = min(re_, int(self.offsets[bisect.bisect_left(self.users, pf.user_lt)]))
        re_ = max(re_, rb)
        mask = None
        if pf.category_values is not None:
            mask = self._mask_of(pf.category_values)
        if pf.category_excluded:
            if self.has_missing_category:
                raise NotImplementedError("must_not on metadata.category needs every row to carry a category "
                                          "(rows without one must pass, and the category mask cannot say so)")
            excl = self._mask_of(pf.category_excluded)
            if excl:
                mask = (_lib.FCB_ALL_CATEGORIES if mask is None else mask) & ~excl
        date_lo = pf.date_lo
        if self.has_missing_date and (pf.date_lo is not None or pf.date_hi is not None):
            # rows without metadata.date sit at INT32_MIN: Qdrant's Range never matches a missing key (and must_not on it
            # passes). A lower bound above the sentinel keeps them out; an upper bound would let them in -> refuse.
            if pf.date_hi is not None:
                raise NotImplementedError("Range(lte/lt) and must_not on metadata.date need every row to carry a date "
                                          "(this corpus has rows without one)")
            date_lo = max(int(pf.date_lo), -(2 ** 31) + 1)
        return make_filter(rb, re_, date_lo, pf.date_hi, mask)

    def resolve_dnf(self, flt: Any) -> list:
        """qdrant filter -> the FcbFilters whose union it selects (empty conjuncts dropped; may be [])."""
        out = []
        for pf in parse_filter_dnf(flt):
            f = self.resolve(pf)
            if f.row_begin < f.row_end and f.date_lo <= f.date_hi and f.category_mask != 0:
                out.append(f)
        return out

    def search_union(self, search: Callable, q: np.ndarray, filters: list, k: int):
        """Exact top-k over the union of ``filters`` for ONE query: all conjuncts in one grouped call
        (chunked by the engine's batch limit). ``search(queries[B, dim], filters, k) -> ids, scores, counts``.
        -> (ids[n], scores[n]) with n <= k."""
        if not filters:
            return np.zeros(0, dtype=np.uint64), np.zeros(0, dtype=np.float64)
        if len(filters) == 1:
            ids, sc, cn = search(q[None, :], filters, k)
            n = int(cn[0])
            return ids[0, :n], sc[0, :n]
        parts = []
        step = max(1, int(getattr(self.engine, "max_batch", len(filters))))
        for lo in range(0, len(filters), step):
            fl = filters[lo:lo + step]
            parts.append(search(np.repeat(q[None, :], len(fl), axis=0), fl, k))
        ids = np.concatenate([p[0] for p in parts]); sc = np.concatenate([p[1] for p in parts])
        cn = np.concatenate([p[2] for p in parts])
        return merge_conjunct_lists(ids, sc, cn, k)


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
        ids, scores = col.search_union(col.engine.search_host, q, col.resolve_dnf(query_filter), k)
        points: List[ScoredPoint] = []
        for j in range(skip, ids.shape[0]):
            s = float(scores[j])
            if score_threshold is not None and s < score_threshold:
                break
            pid = int(ids[j])
            row = col.row_of_id(pid)
            payload = col.payload_fn(row) if with_payload else None
            points.append(ScoredPoint(id=col.point_id_fn(row) if col.point_id_fn else pid, version=0, score=s,
                                      payload=payload, vector=None))
        return QueryResponse(points=points)
