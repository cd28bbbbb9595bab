"""Host-side mirror of the reference interface: filter resolution, query_points plumbing,
retrieve_transactions behaviour (reference tools/qdrant_tool.py:75-177). No GPU: the engine is
replaced by a recording fake so only the host logic is exercised."""
import datetime as dt

import numpy as np
import pytest

from financial_chatbot_llm_b200 import _lib, client, tools
from financial_chatbot_llm_b200 import qdrant_models as m


def test_range_bounds_are_inclusive_integers():
    pf = client.parse_filter(m.Filter(must=[
        m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="u1")),
        m.FieldCondition(key="metadata.date", range=m.Range(gte=10.5, lt=20)),
        m.FieldCondition(key="metadata.date", range=m.Range(gt=11, lte=30.9))]))
    assert pf.has_user and pf.user_id == "u1" and (pf.date_lo, pf.date_hi) == (12, 19)
    assert client.parse_filter(None).has_user is False


def test_unsupported_filters_raise():
    with pytest.raises(NotImplementedError):
        client.parse_filter(m.Filter(min_should=2, should=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="a"))]))
    with pytest.raises(NotImplementedError):
        client.parse_filter(m.Filter(must=[m.FieldCondition(key="metadata.amount", range=m.Range(gte=1))]))
    with pytest.raises(NotImplementedError):                                   # two conjuncts do not fit ONE conjunction
        client.parse_filter(m.Filter(must_not=[m.FieldCondition(key="metadata.date", range=m.Range(gte=5, lte=9))]))
    # a single should is just a must; a filter nothing can pass resolves to an empty segment
    pf = client.parse_filter(m.Filter(should=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="a"))]))
    assert pf.has_user and pf.user_id == "a"
    pf = client.parse_filter(m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="a")),
                                            m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="b"))]))
    assert pf.has_user and pf.user_id is None


def test_should_and_must_not_rewrite_selects_exactly_what_the_filter_model_selects():
    """should / must_not become a union of conjunctions (row range AND date range AND category mask):
    brute-force the union over a small table and compare with the oracle's per-payload filter check."""
    from oracle import local_mode
    rng = np.random.default_rng(7)
    users = ["u%02d" % i for i in range(6)]
    cats = ["Rent", "Fuel", "Dining", "Other"]
    sizes = [5, 0, 9, 1, 7, 4]                                   # one user owns no rows
    offsets = np.concatenate([[0], np.cumsum(sizes)])
    n = int(offsets[-1])
    row_user = np.repeat(np.arange(6), sizes)
    dates = rng.integers(0, 30, n)
    codes = rng.integers(0, 4, n)
    payloads = [{"page_content": "x", "metadata": {"user_id": users[row_user[r]], "date": int(dates[r]),
                                                   "category": cats[codes[r]]}} for r in range(n)]
    col = client.Collection("t", FakeEngine(), users, offsets, lambda r: payloads[r],
                            category_codes={c: i for i, c in enumerate(cats)})

    def U(v): return m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=v))
    def D(**kw): return m.FieldCondition(key="metadata.date", range=m.Range(**kw))
    def C(*v): return m.FieldCondition(key="metadata.category", match=m.MatchAny(any=list(v)))
    filters = [
        m.Filter(must_not=[U("u02")]),
        m.Filter(must_not=[U("u02"), U("u04"), U("zz")]),
        m.Filter(must=[U("u02")], must_not=[D(gte=10, lte=20)]),
        m.Filter(must_not=[D(gte=10)]), m.Filter(must_not=[D(lt=12.5)]), m.Filter(must_not=[D()]),
        m.Filter(must_not=[C("Rent", "Fuel")]), m.Filter(must=[C("Rent", "Dining")], must_not=[C("Rent")]),
        m.Filter(should=[U("u00"), U("u05"), U("u01")]),
        m.Filter(must=[D(gte=5)], should=[C("Rent"), D(lte=8), U("u04")], must_not=[U("u02"), D(gte=15, lt=18)]),
        m.Filter(must=[U("u02")], must_not=[U("u02")]), m.Filter(should=[]), m.Filter(must_not=[C("Nope")]),
        m.Filter(must=[U("u04")], should=[D(lte=3), D(gte=25)], must_not=[C("Other")]),
    ]
    rows = np.arange(n)
    for flt in filters:
        got = np.zeros(n, dtype=bool)
        for f in col.resolve_dnf(flt):
            got |= (rows >= f.row_begin) & (rows < f.row_end) & (dates >= f.date_lo) & (dates <= f.date_hi) & \
                   (((f.category_mask >> codes.astype(np.uint64)) & 1) == 1)
        want = local_mode.payload_mask(payloads, flt)
        assert np.array_equal(got, want), flt
    # rows without a category must pass must_not(category): the mask cannot say that, so it is refused
    col.has_missing_category = True
    with pytest.raises(NotImplementedError):
        col.resolve_dnf(m.Filter(must_not=[C("Rent")]))


def test_union_of_conjunct_lists_drops_duplicates_and_keeps_the_order():
    ids = np.array([[5, 9, 2, 0], [9, 7, 0, 0], [1, 0, 0, 0]], dtype=np.uint64)
    sc = np.array([[0.9, 0.8, 0.8, 0], [0.8, 0.85, 0, 0], [0.8, 0, 0, 0]])
    i, s = client.merge_conjunct_lists(ids, sc, np.array([3, 2, 1]), 4)
    assert i.tolist() == [5, 7, 1, 2] and s.tolist() == [0.9, 0.85, 0.8, 0.8]


class FakeEngine:
    dim, row_base = 4, 0

    def __init__(self):
        self.calls = []

    def search_host(self, q, filters, k):
        self.calls.append((q.copy(), filters[0], k))
        f = filters[0]
        rows = np.arange(f.row_begin, f.row_end, dtype=np.uint64)[:k]
        ids = np.full((1, k), np.iinfo(np.uint64).max, dtype=np.uint64)
        sc = np.full((1, k), -np.inf)
        ids[0, :rows.size] = rows
        sc[0, :rows.size] = 1.0 - 0.01 * np.arange(rows.size)
        return ids, sc, np.array([rows.size], dtype=np.int32)


def make_client(wrong_user_rows=()):
    eng = FakeEngine()
    users = ["alice", "bob"]
    offsets = np.array([0, 3, 8])

    def payload(row):
        u = "mallory" if row in wrong_user_rows else users[0 if row < 3 else 1]
        return {"page_content": f"txn {row}", "metadata": {"user_id": u, "date": 100 + row}}
    col = client.Collection("transactions", eng, users, offsets, payload, category_codes={"Rent": 2})
    c = client.B200QdrantClient(url="http://ignored", api_key="ignored")
    c.attach_collection(col)
    return c, eng


def test_query_points_shape_and_filter_resolution():
    c, eng = make_client()
    flt = m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="bob")),
                         m.FieldCondition(key="metadata.category", match=m.MatchValue(value="Rent"))])
    res = c.query_points(collection_name="transactions", query=[1, 0, 0, 0], limit=3,
                         search_params=m.SearchParams(hnsw_ef=128, exact=False), query_filter=flt)
    assert [p.id for p in res.points] == [3, 4, 5] and res.points[0].version == 0
    assert res.points[0].payload["page_content"] == "txn 3" and res.points[0].vector is None
    f = eng.calls[0][1]
    assert (f.row_begin, f.row_end, f.category_mask) == (3, 8, 1 << 2)
    assert f.date_lo == _lib.FCB_DATE_MIN and f.date_hi == _lib.FCB_DATE_MAX
    # unknown user -> empty segment; offset and score_threshold are applied host-side
    assert c.query_points("transactions", query=[1, 0, 0, 0], limit=5, query_filter=m.Filter(must=[
        m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="nobody"))])).points == []
    pts = c.query_points("transactions", query=[1, 0, 0, 0], limit=2, offset=1, score_threshold=0.985,
                         query_filter=flt).points
    assert [p.id for p in pts] == [4]
    with pytest.raises(ValueError):
        c.query_points("missing", query=[1, 0, 0, 0])


def test_retrieve_transactions_behaviour():
    c, eng = make_client(wrong_user_rows={4})
    now = dt.datetime(2024, 6, 1, 12, 0, 0)
    tools.configure(c, lambda text: [0.0, 1.0, 0.0, 0.0], now=lambda: now)
    assert tools.retrieve_transactions(user_id="") == []                       # :89-91, no search
    assert eng.calls == []
    out = tools.retrieve_transactions(user_id="bob", num_transactions=4, time_period_days=30)
    assert out == ["txn 3", "txn 5", "txn 6"]                                  # row 4 fails the post-check
    _, f, k = eng.calls[-1]
    assert k == 4 and f.date_lo == int((now - dt.timedelta(days=30)).timestamp())
    tools.retrieve_transactions(user_id="bob", time_period_days=0)             # 0 => no date filter (:116)
    _, f, k = eng.calls[-1]
    assert k == 10000 and f.date_lo == _lib.FCB_DATE_MIN                       # default limit (:145)
    tools.configure(c, lambda text: 1 / 0)                                     # any exception -> [] (:175-177)
    assert tools.retrieve_transactions(user_id="bob") == []
    with pytest.raises(Exception):
        tools.RetrievalIntent(num_transactions=10001)


def test_random_filters_rewrite_equals_the_filter_model():
    """300 seeded random must / should / must_not combinations over the three keys (including unknown users, unknown
    categories, empty and inverted ranges, fractional bounds): the union of the rewritten conjunctions selects exactly the
    rows the oracle's per-payload check selects, or the rewrite refuses (too many conjunctions) — never a wrong set."""
    from oracle import local_mode
    rng = np.random.default_rng(2024)
    users = ["u%02d" % i for i in range(8)]
    cats = ["Rent", "Fuel", "Dining", "Other", "Travel"]
    sizes = [4, 0, 7, 1, 9, 3, 0, 6]
    offsets = np.concatenate([[0], np.cumsum(sizes)])
    n = int(offsets[-1])
    row_user = np.repeat(np.arange(8), sizes)
    dates = rng.integers(-5, 40, n)
    codes = rng.integers(0, 5, n)
    payloads = [{"page_content": "x", "metadata": {"user_id": users[row_user[r]], "date": int(dates[r]),
                                                   "category": cats[codes[r]]}} for r in range(n)]
    col = client.Collection("t", FakeEngine(), users, offsets, lambda r: payloads[r],
                            category_codes={c: i for i, c in enumerate(cats)})

    def cond():
        kind = rng.integers(0, 3)
        if kind == 0:
            return m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=str(rng.choice(users + ["zz", "a"]))))
        if kind == 1:
            kw = {}
            for name in ("gte", "gt", "lte", "lt"):
                if rng.random() < 0.4:
                    kw[name] = float(rng.integers(-8, 45)) + float(rng.choice([0.0, 0.5]))
            return m.FieldCondition(key="metadata.date", range=m.Range(**kw))
        pick = [str(c) for c in rng.choice(cats + ["Nope"], size=int(rng.integers(1, 4)), replace=False)]
        return m.FieldCondition(key="metadata.category", match=m.MatchAny(any=pick) if len(pick) > 1 or rng.random() < 0.5
                                else m.MatchValue(value=pick[0]))
    rows = np.arange(n)
    checked = refused = 0
    for _ in range(300):
        flt = m.Filter(must=[cond() for _ in range(rng.integers(0, 3))] or None,
                       should=[cond() for _ in range(rng.integers(0, 4))] or None,
                       must_not=[cond() for _ in range(rng.integers(0, 4))] or None)
        try:
            conj = col.resolve_dnf(flt)
        except NotImplementedError:
            refused += 1
            continue
        got = np.zeros(n, dtype=bool)
        for f in conj:
            got |= (rows >= f.row_begin) & (rows < f.row_end) & (dates >= f.date_lo) & (dates <= f.date_hi) & \
                   (((f.category_mask >> codes.astype(np.uint64)) & 1) == 1)
        assert np.array_equal(got, local_mode.payload_mask(payloads, flt)), flt
        checked += 1
    assert checked > 250 and refused < 50
