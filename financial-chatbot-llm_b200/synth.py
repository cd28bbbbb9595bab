"""Seeded synthetic transactions corpus (host side, numpy).

The reference has no ingestion code and no data, so every
test and bench line runs on synthetic points. The generator is COUNTER-BASED
and INTEGER-ONLY: every element is a pure function of (seed, stream, index)
through a splitmix64 finaliser, converted to float by an exact power-of-two
scaling. The same recipe is implemented on the device in csrc/synth.cu, so a
10M-row corpus can be generated straight into HBM and any row range of it can
be regenerated bit-identically on the host for a cross-check (no libm calls,
no reductions, nothing whose rounding could differ between CPU and GPU).

Layout of one point (mirrors the payload shape read by the reference at
tools/qdrant_tool.py:107,120,160-164 — ``metadata.user_id``, ``metadata.date``
unix seconds, ``page_content``):

  vector    d floats: cluster centre (one of ``n_clusters``) + ~Gaussian noise
  user      CSR segment index (rows are sorted by user)
  date      int32 unix seconds, uniform over ``date_span_days``
  category  uint8 in [0, n_categories)
  cents     int32 signed integer cents, log-uniform magnitude
  id        uint64 (row index unless a permutation is requested)
"""
from __future__ import annotations

import dataclasses
from typing import Optional, Sequence

import numpy as np

U64 = np.uint64
_GOLDEN = U64(0x9E3779B97F4A7C15)
_STREAM_MUL = U64(0xD6E8FEB86659FD93)
_M1 = U64(0xBF58476D1CE4E5B9)
_M2 = U64(0x94D049BB133111EB)

# stream ids (shared with csrc/synth.cu)
S_CLUSTER, S_CENTRE, S_NOISE, S_DATE, S_CAT, S_CENTS, S_QUERY, S_QCLUSTER = range(1, 9)

DATE_BASE = 1_640_995_200          # 2022-01-01T00:00:00Z
VALUE_SCALE = 2.0 ** -18           # exact power-of-two scaling of the integer element
CATEGORIES = ("Groceries", "Dining", "Rent", "Utilities", "Transport", "Fuel",
              "Entertainment", "Travel", "Healthcare", "Insurance", "Shopping",
              "Subscriptions", "Education", "Income", "Transfers", "Other")


def mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on a uint64 array (wrapping arithmetic)."""
    x = x.astype(U64, copy=True)
    x ^= x >> U64(30)
    x *= _M1
    x ^= x >> U64(27)
    x *= _M2
    x ^= x >> U64(31)
    return x


def hash_u64(seed: int, stream: int, idx: np.ndarray) -> np.ndarray:
    """h(seed, stream, idx) = mix64((seed ^ stream*K) + idx*GOLDEN)."""
    with np.errstate(over="ignore"):
        base = U64(seed & 0xFFFFFFFFFFFFFFFF) ^ (U64(stream) * _STREAM_MUL)
        return mix64(base + np.asarray(idx, dtype=U64) * _GOLDEN)


def _noise_from_bits(u: np.ndarray) -> np.ndarray:
    """Sum of four 16-bit uniforms, centred: integer in [-131070, 131070]."""
    m = U64(0xFFFF)
    s = (u & m) + ((u >> U64(16)) & m) + ((u >> U64(32)) & m) + (u >> U64(48))
    return s.astype(np.int64) - 131070


@dataclasses.dataclass(frozen=True)
class SynthSpec:
    n_rows: int
    dim: int = 1536
    seed: int = 20240607
    n_clusters: int = 256
    n_categories: int = 16
    date_span_days: int = 730
    n_users: int = 1000
    zipf_a: float = 1.1            # segment-size skew; 0 => equal sizes
    whale_rows: int = 0            # >0: user 0 owns exactly this many rows, rest Zipf

    @property
    def date_span_seconds(self) -> int:
        return self.date_span_days * 86400


def segment_offsets(spec: SynthSpec) -> np.ndarray:
    """CSR offsets (int64, n_users+1) for rows sorted by user. Deterministic,
    pure integer after the float weights are floored; every user gets >= 1 row."""
    n, u = spec.n_rows, spec.n_users
    if u <= 0 or n < u:
        raise ValueError("need 1 <= n_users <= n_rows")
    sizes = np.ones(u, dtype=np.int64)
    first = 0
    remaining = n - u
    if spec.whale_rows:
        if spec.whale_rows > n - (u - 1):
            raise ValueError("whale_rows leaves no rows for the other users")
        sizes[0] = spec.whale_rows
        remaining = n - spec.whale_rows - (u - 1)
        first = 1
    m = u - first
    if m > 0 and remaining > 0:
        w = 1.0 / np.power(np.arange(1, m + 1, dtype=np.float64), spec.zipf_a)
        extra = np.floor(w / w.sum() * remaining).astype(np.int64)
        extra[0] += remaining - int(extra.sum())
        sizes[first:] += extra
    off = np.zeros(u + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    assert off[-1] == n
    return off


def user_ids(n_users: int) -> list[str]:
    """Fixed-width ids so lexicographic order == segment order."""
    return [f"user_{i:07d}" for i in range(n_users)]


def row_clusters(spec: SynthSpec, rows: np.ndarray) -> np.ndarray:
    return (hash_u64(spec.seed, S_CLUSTER, rows) % U64(spec.n_clusters)).astype(np.int64)


def centre_ints(spec: SynthSpec, clusters: np.ndarray) -> np.ndarray:
    """int64 [len(clusters), dim] centre components in [-32768, 32767]."""
    c = np.asarray(clusters, dtype=U64)[:, None]
    j = np.arange(spec.dim, dtype=U64)[None, :]
    h = hash_u64(spec.seed, S_CENTRE, c * U64(spec.dim) + j)
    return (h >> U64(48)).astype(np.int64) - 32768


def vectors_int(spec: SynthSpec, row_lo: int, row_hi: int) -> np.ndarray:
    """Exact integer elements (int64 [rows, dim]); |v| < 2**24 so float32 holds them exactly."""
    rows = np.arange(row_lo, row_hi, dtype=U64)
    cl = row_clusters(spec, rows)
    j = np.arange(spec.dim, dtype=U64)[None, :]
    u = hash_u64(spec.seed, S_NOISE, rows[:, None] * U64(spec.dim) + j)
    return 4 * centre_ints(spec, cl) + _noise_from_bits(u)


def vectors_f32(spec: SynthSpec, row_lo: int, row_hi: int) -> np.ndarray:
    return (vectors_int(spec, row_lo, row_hi).astype(np.float32) * np.float32(VALUE_SCALE))


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even float32 -> bfloat16 bit pattern (uint16). Finite inputs only."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    r = (u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) >> np.uint32(16)
    return r.astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.asarray(b, dtype=np.uint16).astype(np.uint32) << np.uint32(16)).view(np.float32)


def vectors_bf16_bits(spec: SynthSpec, row_lo: int, row_hi: int) -> np.ndarray:
    return f32_to_bf16_bits(vectors_f32(spec, row_lo, row_hi))


def dates(spec: SynthSpec, row_lo: int, row_hi: int) -> np.ndarray:
    rows = np.arange(row_lo, row_hi, dtype=U64)
    h = hash_u64(spec.seed, S_DATE, rows)
    return (np.int64(DATE_BASE) + (h % U64(spec.date_span_seconds)).astype(np.int64)).astype(np.int32)


def dates_sorted(spec: SynthSpec, offsets: np.ndarray, row_lo: int, row_hi: int) -> np.ndarray:
    """Time-ordered variant: row r of user segment [a, b) gets DATE_BASE + floor((r - a) * span / (b - a))."""
    rows = np.arange(row_lo, row_hi, dtype=np.int64)
    u = row_users(offsets, rows)
    a, b = offsets[u], offsets[u + 1]
    return (np.int64(DATE_BASE) + (rows - a) * np.int64(spec.date_span_seconds) // (b - a)).astype(np.int32)


def categories(spec: SynthSpec, row_lo: int, row_hi: int) -> np.ndarray:
    rows = np.arange(row_lo, row_hi, dtype=U64)
    return (hash_u64(spec.seed, S_CAT, rows) % U64(spec.n_categories)).astype(np.uint8)


def cents(spec: SynthSpec, row_lo: int, row_hi: int) -> np.ndarray:
    """Signed integer cents: magnitude log-uniform over ~5 decades, 1/8 negative."""
    rows = np.arange(row_lo, row_hi, dtype=U64)
    u = hash_u64(spec.seed, S_CENTS, rows)
    e = (u & U64(0xFF)) % U64(18)
    mant = U64(1024) + ((u >> U64(8)) & U64(1023))
    mag = ((mant << e) >> U64(10)).astype(np.int64)
    neg = ((u >> U64(40)) & U64(7)) == U64(0)
    return np.where(neg, -mag, mag).astype(np.int32)


def query_vector(spec: SynthSpec, qidx: int) -> np.ndarray:
    """float64 query #qidx: a cluster centre + independent noise (what embed_query
    would hand to query_points as list[float], tools/qdrant_tool.py:137)."""
    q = np.array([qidx], dtype=U64)
    cl = (hash_u64(spec.seed, S_QCLUSTER, q) % U64(spec.n_clusters)).astype(np.int64)
    j = np.arange(spec.dim, dtype=U64)[None, :]
    u = hash_u64(spec.seed, S_QUERY, q[:, None] * U64(spec.dim) + j)
    vi = 4 * centre_ints(spec, cl) + _noise_from_bits(u)
    return (vi[0].astype(np.float64) * VALUE_SCALE)


def row_users(offsets: np.ndarray, rows: np.ndarray) -> np.ndarray:
    return np.searchsorted(offsets, rows, side="right") - 1


def page_content(user: str, date: int, category: int, cents_v: int, row: int) -> str:
    """Text a LangChain-style ingester would have stored as page_content."""
    import datetime as _dt
    day = _dt.datetime.fromtimestamp(int(date), _dt.timezone.utc).strftime("%Y-%m-%d")
    sign = "-" if cents_v < 0 else ""
    a = abs(int(cents_v))
    return (f"{day} {CATEGORIES[int(category) % len(CATEGORIES)]} "
            f"{sign}${a // 100}.{a % 100:02d} merchant_{row % 9973:04d} (txn {row})")


_MASK64 = (1 << 64) - 1


def _hash_int(seed: int, stream: int, idx: int) -> int:
    """hash_u64 for ONE index in plain Python integers (payloads of the k winners: no numpy overhead)."""
    x = (((seed & _MASK64) ^ ((stream * int(_STREAM_MUL)) & _MASK64)) + idx * int(_GOLDEN)) & _MASK64
    x ^= x >> 30
    x = (x * int(_M1)) & _MASK64
    x ^= x >> 27
    x = (x * int(_M2)) & _MASK64
    x ^= x >> 31
    return x


def payload_for_row(spec: SynthSpec, offsets: np.ndarray, uids: Sequence[str], row: int) -> dict:
    """Payload dict in the layout the reference reads (tools/qdrant_tool.py:160-164). Same values as the
    column generators (dates / categories / cents) for this row."""
    u = int(np.searchsorted(offsets, row, side="right")) - 1
    d = DATE_BASE + _hash_int(spec.seed, S_DATE, row) % spec.date_span_seconds
    c = _hash_int(spec.seed, S_CAT, row) % spec.n_categories
    h = _hash_int(spec.seed, S_CENTS, row)
    mag = ((1024 + ((h >> 8) & 1023)) << ((h & 0xFF) % 18)) >> 10
    v = -mag if ((h >> 40) & 7) == 0 else mag
    return {
        "page_content": page_content(uids[u], d, c, v, row),
        "metadata": {"user_id": uids[u], "date": d, "category": CATEGORIES[c % len(CATEGORIES)],
                     "amount_cents": v},
    }


@dataclasses.dataclass
class HostCorpus:
    """A fully materialised small corpus (tests, config 1)."""
    spec: SynthSpec
    vectors: np.ndarray            # float32 [n, d] (bf16 corpora: the bf16-rounded values)
    ids: np.ndarray                # uint64 [n]
    offsets: np.ndarray            # int64 [n_users+1]
    users: list[str]
    dates: np.ndarray              # int32 [n]
    categories: np.ndarray         # uint8 [n]
    cents: np.ndarray              # int32 [n]
    bf16_bits: Optional[np.ndarray] = None

    def payload(self, row: int) -> dict:
        u = int(row_users(self.offsets, np.array([row]))[0])
        d, c, v = int(self.dates[row]), int(self.categories[row]), int(self.cents[row])
        return {
            "page_content": page_content(self.users[u], d, c, v, int(self.ids[row])),
            "metadata": {"user_id": self.users[u], "date": d,
                         "category": CATEGORIES[c % len(CATEGORIES)], "amount_cents": v},
        }


def make_host_corpus(spec: SynthSpec, bf16: bool = False, chunk: int = 16384) -> HostCorpus:
    n = spec.n_rows
    vec = np.empty((n, spec.dim), dtype=np.float32)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        vec[lo:hi] = vectors_f32(spec, lo, hi)
    bits = None
    if bf16:
        bits = f32_to_bf16_bits(vec).reshape(n, spec.dim)
        vec = bf16_bits_to_f32(bits).reshape(n, spec.dim)
    return HostCorpus(
        spec=spec, vectors=vec, ids=np.arange(n, dtype=np.uint64),
        offsets=segment_offsets(spec), users=user_ids(spec.n_users),
        dates=dates(spec, 0, n), categories=categories(spec, 0, n), cents=cents(spec, 0, n),
        bf16_bits=bits)
