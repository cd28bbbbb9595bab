"""Pins the oracle: analytic known-answer cases + agreement of its independent restatements.

The reference ships no golden vectors for this path and qdrant-client is not installable
(parity unpinned, see oracle/local_mode.py), so the anchors are (1) cases whose answers follow
from the definition of cosine similarity, (2) numpy restatement == C restatement, (3) the
dict-payload filter == the column filter."""
import numpy as np
import pytest

from financial_chatbot_llm_b200 import qdrant_models as m
from financial_chatbot_llm_b200 import synth
from oracle import canon, local_mode


def test_kat_basis_vectors():
    # rows = scaled unit vectors e_j; query = (4,3,0,...): cos = 4/5, 3/5, 0 ...
    d = 32
    rows = np.zeros((5, d), dtype=np.float32)
    for j in range(5):
        rows[j, j] = 2.0 ** j
    q = np.zeros(d)
    q[0], q[1] = 4.0, 3.0
    s = canon.scores(rows, q)
    assert s.tolist() == [0.8, 0.6, 0.0, 0.0, 0.0]
    ids, sc = canon.search(rows, q, 3)
    assert ids.tolist() == [0, 1, 2] and sc.tolist() == [0.8, 0.6, 0.0]      # tie at 0.0 -> lowest id
    s_np = local_mode.cosine_scores(q, rows.copy(), inplace_renorm=True)
    assert np.allclose(s_np, [0.8, 0.6, 0, 0, 0], atol=1e-7)


def test_kat_identical_and_opposite_rows():
    d = 64
    rng = np.random.default_rng(0)
    v = rng.standard_normal(d).astype(np.float32)
    rows = np.stack([v, -v, 2 * v, np.zeros(d, np.float32)])
    s = canon.scores(rows, v.astype(np.float64))
    assert abs(s[0] - 1) < 1e-15 and abs(s[1] + 1) < 1e-15 and s[2] == s[0] and s[3] == 0.0
    ids, _ = canon.search(rows, v.astype(np.float64), 4, ids=np.array([9, 5, 3, 7], dtype=np.uint64))
    assert ids.tolist() == [3, 9, 7, 5]          # ties (rows 0 and 2) by ascending id


def test_zero_query_scores_zero():
    rows = np.ones((3, 32), dtype=np.float32)
    assert canon.scores(rows, np.zeros(32)).tolist() == [0.0, 0.0, 0.0]


def test_numpy_and_c_restatements_agree(small_corpus):
    hc = small_corpus
    for qi in range(4):
        q = synth.query_vector(hc.spec, qi)
        sc = canon.scores(hc.vectors, q)
        s1 = local_mode.cosine_scores(q, hc.vectors.copy(), inplace_renorm=False)
        s2 = local_mode.cosine_scores(q, hc.vectors.copy(), inplace_renorm=True)
        assert np.abs(sc - s1).max() < 1e-14
        assert np.abs(sc - s2).max() < 5e-7       # local mode's in-place float32 renormalisation
        for u in (0, 5, 39):
            mask = local_mode.column_mask(hc.spec.n_rows, hc.offsets, u, None, None, None, None, None)
            i_np, s_np = local_mode.search_columns(hc.vectors, hc.ids, mask, q, 10)
            i_c, s_c = canon.search(hc.vectors, q, 10, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
            assert i_np.tolist() == i_c.tolist()
            assert np.abs(s_np - s_c).max() < 1e-14


def test_bf16_corpus_is_scored_from_the_rounded_values(small_corpus):
    hc = small_corpus
    bits = synth.f32_to_bf16_bits(hc.vectors).reshape(hc.vectors.shape)
    q = synth.query_vector(hc.spec, 1)
    assert np.array_equal(canon.scores(bits, q), canon.scores(synth.bf16_bits_to_f32(bits).reshape(bits.shape), q))


def _payload_filter(user=None, gte=None, lt=None, cats=None):
    must = []
    if user is not None:
        must.append(m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=user)))
    if gte is not None or lt is not None:
        must.append(m.FieldCondition(key="metadata.date", range=m.Range(gte=gte, lt=lt)))
    if cats is not None:
        must.append(m.FieldCondition(key="metadata.category", match=m.MatchAny(any=cats)))
    return m.Filter(must=must)


def test_dict_payload_filter_equals_column_filter():
    spec = synth.SynthSpec(n_rows=800, dim=32, n_users=9, seed=5)
    hc = synth.make_host_corpus(spec)
    payloads = [hc.payload(r) for r in range(spec.n_rows)]
    t0 = int(np.percentile(hc.dates, 30))
    t1 = int(np.percentile(hc.dates, 80))
    cats = ["Rent", "Fuel", "Other"]
    cmask = sum(1 << synth.CATEGORIES.index(c) for c in cats)
    cases = [
        (_payload_filter(user=hc.users[2]), (2, None, None, None)),
        (_payload_filter(user=hc.users[2], gte=t0), (2, t0, None, None)),
        (_payload_filter(user=hc.users[0], gte=t0, lt=t1), (0, t0, t1 - 1, None)),
        (_payload_filter(user=hc.users[4], cats=cats), (4, None, None, cmask)),
        (_payload_filter(user="nobody"), (-1, None, None, None)),
    ]
    for flt, (u, lo, hi, cm) in cases:
        a = local_mode.payload_mask(payloads, flt)
        b = local_mode.column_mask(spec.n_rows, hc.offsets, u, hc.dates, lo, hi, hc.categories, cm)
        assert np.array_equal(a, b)
    # Range on a non-numeric value and a missing key are False
    assert not local_mode.check_condition(m.FieldCondition(key="metadata.user_id", range=m.Range(gte=0)), payloads[0])
    assert not local_mode.check_condition(m.FieldCondition(key="metadata.nope", match=m.MatchValue(value=1)), payloads[0])


def test_local_collection_query_points_matches_columns():
    spec = synth.SynthSpec(n_rows=600, dim=64, n_users=6, seed=3)
    hc = synth.make_host_corpus(spec)
    col = local_mode.LocalCollection(spec.dim)
    col.upsert([int(i) for i in hc.ids], hc.vectors, [hc.payload(r) for r in range(spec.n_rows)])
    q = synth.query_vector(spec, 0)
    t0 = int(np.median(hc.dates))
    pts = col.query_points(q.tolist(), _payload_filter(user=hc.users[1], gte=t0), limit=7)
    i_c, s_c = canon.search(hc.vectors, q, 7, row_begin=int(hc.offsets[1]), row_end=int(hc.offsets[2]),
                            dates=hc.dates, date_lo=t0)
    assert [p["id"] for p in pts] == i_c.tolist()
    assert np.allclose([p["score"] for p in pts], s_c, rtol=0, atol=5e-7)
    assert all(p["payload"]["metadata"]["user_id"] == hc.users[1] and p["version"] == 0 for p in pts)
    # offset / limit larger than the segment
    many = col.query_points(q.tolist(), _payload_filter(user=hc.users[5]), limit=10000)
    assert len(many) == int(hc.offsets[6] - hc.offsets[5])
    assert col.query_points(q.tolist(), _payload_filter(user="nobody"), limit=10) == []


def test_month_index_against_numpy_calendar():
    from oracle import groupby
    ts = np.array([0, 86399, 86400 * 31, 951782400, 1709164800, 1709251199, 1709251200, -1, -86400 * 31,
                   2**31 - 1, -(2**31)], dtype=np.int64)
    assert [canon.month_index(int(t)) for t in ts] == groupby.month_index(ts).tolist()


def test_groupby_c_restatement_equals_pandas():
    from oracle import groupby
    spec = synth.SynthSpec(n_rows=20000, dim=8, n_users=37, seed=9)
    off = synth.segment_offsets(spec)
    d, c, v = synth.dates(spec, 0, 20000), synth.categories(spec, 0, 20000), synth.cents(spec, 0, 20000)
    users = synth.row_users(off, np.arange(20000))
    ser = groupby.groupby_user_category_month(users, c, d, v)
    u2, c2, m2, s2, _ = canon.groupby(off, d, c, v)
    idx = ser.index.to_frame(index=False)
    assert idx["user"].tolist() == u2.tolist() and idx["category"].tolist() == c2.tolist()
    assert idx["month"].tolist() == m2.tolist() and ser.to_numpy().tolist() == s2.tolist()
    assert ser.dtype == np.int64
