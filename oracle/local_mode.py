"""ORACLE — test infrastructure only. PARITY UNPINNED for retrieval.

CPU restatement (numpy) of the arithmetic behind the reference's hot-path call
``client.query_points(collection_name, query, limit, search_params,
query_filter).points`` (reference tools/qdrant_tool.py:147-153) and of the
filter it builds (tools/qdrant_tool.py:105-131).

The arithmetic itself lives in a THIRD-PARTY dependency that is absent from
/root/reference and from this image: ``qdrant-client==1.14.2``
(reference requirements.txt:78). Its exact in-memory "local mode"
(``QdrantClient(":memory:")``) is the behaviour north_star names as the
correctness bar. The package cannot be imported or installed here (no network),
and the reference ships no tests, golden vectors or fixtures for this path
(SURVEY.md §4), so this restatement follows the published algorithm of
``qdrant_client/local/{local_collection,distances,payload_filters,
payload_value_extractor}.py`` as recalled in SURVEY.md §3.4/§8c and is anchored
on the reference's own call site. => "parity unpinned": nothing here has been
checked against an output of qdrant-client itself. What CAN be checked is:
analytic known-answer cases (tests/test_oracle_kat.py), agreement between this
numpy restatement and the independent C restatement (oracle/canon.c), and the
reference's own ``retrieve_transactions`` body executed over this oracle
(tests/golden/make_golden.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module. The product never does.

Local-mode steps restated:
  1. payload mask: AND over ``must`` conditions, OR over ``should``, NOT over
     ``must_not``; dotted keys resolved into a list of values; MatchValue =
     any value ==; MatchAny = any value in set; Range = numeric and inside
     every non-None bound; missing key -> False.
  2. cosine: stored float32 rows are re-normalised, the float64 query is
     normalised, scores = rows . query over ALL rows (filter does not prune).
  3. order by score descending. numpy's argsort leaves ties unspecified; the
     oracle DEFINES ties as ascending point id (north_star: "ties ... resolved
     by the same id order").
  4. first ``limit`` masked-in points -> (id, version=0, score, payload).
"""
from __future__ import annotations

from typing import Any, Callable, Iterable, List, Optional, Sequence

import numpy as np

EPSILON = 1.1920929e-7      # float32 eps: zero-norm guard used by local mode's distances


# --------------------------------------------------------------------------- payload filter
def _extract(payload: Any, key: str) -> list:
    """Values at a dotted key path; arrays are flattened; missing -> []."""
    cur = [payload]
    for part in key.split("."):
        nxt = []
        for c in cur:
            if isinstance(c, dict) and part in c:
                v = c[part]
                nxt.extend(v) if isinstance(v, list) else nxt.append(v)
        cur = nxt
    return cur


def _check_match(match: Any, values: list) -> bool:
    if hasattr(match, "value"):
        return any(v == match.value for v in values)
    if hasattr(match, "any"):
        allowed = set(match.any)
        return any(v in allowed for v in values)
    raise NotImplementedError(f"match type {type(match).__name__}")


def _check_range(rng: Any, values: list) -> bool:
    def ok(v):
        if isinstance(v, bool) or not isinstance(v, (int, float)):
            return False
        return ((rng.lt is None or v < rng.lt) and (rng.lte is None or v <= rng.lte)
                and (rng.gt is None or v > rng.gt) and (rng.gte is None or v >= rng.gte))
    return any(ok(v) for v in values)


def check_condition(cond: Any, payload: Any) -> bool:
    values = _extract(payload, cond.key)
    if getattr(cond, "match", None) is not None:
        return _check_match(cond.match, values)
    if getattr(cond, "range", None) is not None:
        return _check_range(cond.range, values)
    raise NotImplementedError("condition without match/range")


def check_filter(flt: Any, payload: Any) -> bool:
    if flt is None:
        return True
    must = getattr(flt, "must", None)
    should = getattr(flt, "should", None)
    must_not = getattr(flt, "must_not", None)
    if must and not all(check_condition(c, payload) for c in must):
        return False
    if should and not any(check_condition(c, payload) for c in should):
        return False
    if must_not and any(check_condition(c, payload) for c in must_not):
        return False
    return True


def payload_mask(payloads: Sequence[Any], flt: Any) -> np.ndarray:
    """Step 1, the way local mode does it: an interpreted loop over every payload."""
    return np.fromiter((check_filter(flt, p) for p in payloads), dtype=bool, count=len(payloads))


# --------------------------------------------------------------------------- distance
def cosine_scores(query: np.ndarray, vectors: np.ndarray, *, inplace_renorm: bool = True) -> np.ndarray:
    """Step 2. ``vectors`` float32 [n,d]; ``query`` any float dtype (a python list
    becomes float64, so the dot is mixed f32 x f64 -> float64 scores).

    inplace_renorm=True follows local mode literally: the stored float32 matrix is
    divided by its row norms IN PLACE on every call (so stored rows drift by
    <= 1 ulp per call until they are fixed points of the operation).
    inplace_renorm=False leaves the matrix untouched and divides the float64 dot
    by float64 norms — the same cosine to within 1.2e-7 relative."""
    q = np.array(query)                      # copy; float64 for list input
    qn = np.linalg.norm(q)
    q = q / np.where(qn != 0.0, qn, EPSILON)
    if inplace_renorm:
        norms = np.linalg.norm(vectors, axis=-1)[:, np.newaxis]
        vectors /= np.where(norms != 0.0, norms, EPSILON)
        return np.dot(vectors, q)
    v64n = np.sqrt(np.einsum("ij,ij->i", vectors, vectors, dtype=np.float64))
    return np.dot(vectors, q) / np.where(v64n != 0.0, v64n, EPSILON)


def order_desc_ties_by_id(scores: np.ndarray, ids: np.ndarray) -> np.ndarray:
    """Step 3 with the oracle-defined tie rule: (score desc, id asc)."""
    return np.lexsort((ids, -scores))


# --------------------------------------------------------------------------- collection
class LocalCollection:
    """Dict-payload collection: the slow, literal model of local mode. Tiny N only."""

    def __init__(self, dim: int):
        self.dim = dim
        self.vectors = np.zeros((0, dim), dtype=np.float32)
        self.ids: list = []
        self.payloads: list = []

    def upsert(self, ids: Iterable, vectors: np.ndarray, payloads: Sequence[Any]) -> None:
        """Cosine collections normalise each vector at insert and store float32."""
        v = np.array(vectors, dtype=np.float64)
        n = np.linalg.norm(v, axis=1, keepdims=True)
        v = np.where(n > 0, v / np.where(n > 0, n, 1.0), v)
        self.vectors = np.vstack([self.vectors, v.astype(np.float32)])
        self.ids.extend(ids)
        self.payloads.extend(payloads)

    def query_points(self, query, query_filter=None, limit: int = 10, offset: int = 0,
                     with_payload: bool = True, score_threshold: Optional[float] = None,
                     search_params: Any = None) -> List[dict]:
        # search_params (hnsw_ef / exact) only steer the remote HNSW server; local
        # mode is always exact and ignores them (tools/qdrant_tool.py:98-101).
        mask = payload_mask(self.payloads, query_filter)
        scores = cosine_scores(query, self.vectors, inplace_renorm=True)
        ids = np.asarray(self.ids, dtype=np.uint64)
        out = []
        skip = offset or 0
        for i in order_desc_ties_by_id(scores, ids):
            if not mask[i]:
                continue
            if score_threshold is not None and scores[i] < score_threshold:
                break
            if skip:
                skip -= 1
                continue
            if len(out) >= limit:
                break
            out.append({"id": int(ids[i]), "version": 0, "score": float(scores[i]),
                        "payload": self.payloads[i] if with_payload else None, "vector": None})
        return out


# --------------------------------------------------------------------------- columnar variant
def column_mask(n: int, offsets: np.ndarray, segment: int, dates: Optional[np.ndarray],
                date_lo: Optional[int], date_hi: Optional[int],
                categories: Optional[np.ndarray], category_mask: Optional[int]) -> np.ndarray:
    """Vectorised step 1 for the column layout: the SAME predicate as
    payload_mask() for Filter(must=[user_id MatchValue, date Range, category Match*])
    when users/dates/categories are stored as columns (tests cross-check the two)."""
    m = np.zeros(n, dtype=bool)
    if segment < 0:
        return m
    m[int(offsets[segment]):int(offsets[segment + 1])] = True
    if dates is not None and date_lo is not None:
        m &= dates.astype(np.int64) >= date_lo
    if dates is not None and date_hi is not None:
        m &= dates.astype(np.int64) <= date_hi
    if categories is not None and category_mask is not None:
        lut = np.array([(category_mask >> c) & 1 for c in range(64)], dtype=bool)
        m &= lut[np.minimum(categories, 63)] & (categories < 64)
    return m


def search_columns(vectors: np.ndarray, ids: np.ndarray, mask: np.ndarray, query, limit: int,
                   *, inplace_renorm: bool = False):
    """Steps 2-4 over a column corpus. Returns (ids uint64[k'], scores float64[k'])."""
    scores = cosine_scores(query, vectors, inplace_renorm=inplace_renorm)
    sel = np.flatnonzero(mask)
    if sel.size == 0:
        return np.zeros(0, np.uint64), np.zeros(0, np.float64)
    s, i = scores[sel], ids[sel]
    k = min(limit, sel.size)
    if k < sel.size:
        # partial select first (what a timing-honest local mode would need), then exact order
        kth = np.partition(s, sel.size - k)[sel.size - k]
        keep = s >= kth
        s, i = s[keep], i[keep]
    o = order_desc_ties_by_id(s, i)[:k]
    return i[o].astype(np.uint64), s[o]
