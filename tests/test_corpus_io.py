"""On-disk corpus format: write points in the payload shape the reference reads, map them back, and (GPU)
load them into an engine and search — results equal to the oracle over the same points."""
import json
import os

import numpy as np
import pytest


def make_points(n=900, dim=64, seed=5, uuid_ids=False):
    from financial_chatbot_llm_b200 import synth
    spec = synth.SynthSpec(n_rows=n, dim=dim, n_users=6, seed=seed)
    hc = synth.make_host_corpus(spec)
    rng = np.random.default_rng(seed)
    order = rng.permutation(n)                                  # points arrive in arbitrary order
    pts = []
    for r in order:
        pl = hc.payload(int(r))
        pid = f"00000000-0000-4000-8000-{int(r):012x}" if uuid_ids else int(r) * 7 + 3
        pts.append({"id": pid, "vector": hc.vectors[r].tolist(), "payload": pl})
    return spec, hc, pts


def test_write_and_map_roundtrip(tmp_path):
    from financial_chatbot_llm_b200 import corpus_io, synth
    spec, hc, pts = make_points()
    man = corpus_io.write_corpus(str(tmp_path / "c"), pts, dim=spec.dim, categories=list(synth.CATEGORIES))
    cf = corpus_io.CorpusFiles(str(tmp_path / "c"))
    assert man["n_rows"] == 900 and cf.users == sorted(hc.users) and cf.offsets[-1] == 900
    by_id = {p["id"]: p for p in pts}
    ids = np.array(cf.columns["ids.u64"])
    assert not man["ids_ascending"] and "id_rank.u32" in cf.columns
    assert np.array_equal(np.argsort(np.argsort(ids)), np.array(cf.columns["id_rank.u32"]))
    for row in (0, 17, 450, 899):
        p = by_id[int(ids[row])]
        u = int(np.searchsorted(cf.offsets, row, side="right") - 1)
        assert cf.payload(row) == p["payload"] and p["payload"]["metadata"]["user_id"] == cf.users[u]
        assert np.array_equal(cf.vectors[row], np.asarray(p["vector"], dtype=np.float32))
        assert cf.columns["dates.i32"][row] == p["payload"]["metadata"]["date"]
        assert cf.columns["cents.i32"][row] == p["payload"]["metadata"]["amount_cents"]
        assert synth.CATEGORIES[cf.columns["categories.u8"][row]] == p["payload"]["metadata"]["category"]
    # bf16 storage and UUID ids
    spec2, hc2, pts2 = make_points(n=50, uuid_ids=True)
    man2 = corpus_io.write_corpus(str(tmp_path / "u"), pts2, dim=spec2.dim, dtype="bfloat16")
    cf2 = corpus_io.CorpusFiles(str(tmp_path / "u"))
    assert man2["original_ids"] and cf2.point_ids is not None and cf2.vectors.dtype == np.uint16
    with pytest.raises(ValueError):
        corpus_io.write_corpus(str(tmp_path / "bad"), [{"id": 1, "vector": [0.0] * 3, "payload": pts[0]["payload"]}], dim=64)


def test_rows_without_a_category_get_their_own_code(tmp_path):
    """A payload without metadata.category must not alias the first category: code 255 matches no category filter
    (the kernels test c < 64), and the manifest says so — must_not on the category is then refused, never wrong."""
    from financial_chatbot_llm_b200 import client, corpus_io
    from financial_chatbot_llm_b200 import qdrant_models as m
    spec, hc, pts = make_points(n=40, dim=8)
    for p in pts[::5]:
        del p["payload"]["metadata"]["category"]
    man = corpus_io.write_corpus(str(tmp_path / "m"), pts, dim=spec.dim)
    cf = corpus_io.CorpusFiles(str(tmp_path / "m"))
    cats = np.array(cf.columns["categories.u8"])
    assert man["missing_category"] and int((cats == corpus_io.NO_CATEGORY).sum()) == 8
    assert all(("category" in cf.payload(r)["metadata"]) == (cats[r] != corpus_io.NO_CATEGORY) for r in range(40))
    assert cats[cats != corpus_io.NO_CATEGORY].max() < len(man["categories"])

    class NoEngine:
        dim, row_base, max_batch = 8, 0, 4
    col = client.Collection("t", NoEngine(), cf.users, cf.offsets, cf.payload,
                            category_codes={c: i for i, c in enumerate(man["categories"])})
    col.has_missing_category = bool(man["missing_category"])           # what load_collection sets
    some = man["categories"][0]
    f = col.resolve_dnf(m.Filter(must=[m.FieldCondition(key="metadata.category", match=m.MatchValue(value=some))]))
    assert len(f) == 1 and f[0].category_mask == 1                      # a positive filter is fine: code 255 never passes it
    with pytest.raises(NotImplementedError):
        col.resolve_dnf(m.Filter(must_not=[m.FieldCondition(key="metadata.category", match=m.MatchValue(value=some))]))


def test_rows_without_a_date_never_match_a_range(tmp_path):
    """A payload without metadata.date: Qdrant's Range never matches a missing key. Such rows are stored at INT32_MIN
    (corpus_io.NO_DATE), lower bounds stay above it, and upper bounds / must_not on the date are refused — checked
    against the oracle's per-payload filter."""
    from financial_chatbot_llm_b200 import client, corpus_io
    from financial_chatbot_llm_b200 import qdrant_models as m
    from oracle import local_mode
    spec, hc, pts = make_points(n=40, dim=8)
    for p in pts[::4]:
        del p["payload"]["metadata"]["date"]
    man = corpus_io.write_corpus(str(tmp_path / "d"), pts, dim=spec.dim)
    cf = corpus_io.CorpusFiles(str(tmp_path / "d"))
    dates = np.array(cf.columns["dates.i32"])
    assert man["missing_date"] and int((dates == corpus_io.NO_DATE).sum()) == 10

    class NoEngine:
        dim, row_base, max_batch = 8, 0, 4
    col = client.Collection("t", NoEngine(), cf.users, cf.offsets, cf.payload)
    col.has_missing_date = bool(man["missing_date"])                    # what load_collection sets
    payloads = [cf.payload(r) for r in range(40)]
    for gte in (-(2 ** 40), -(2 ** 31), int(np.median(dates[dates != corpus_io.NO_DATE]))):
        flt = m.Filter(must=[m.FieldCondition(key="metadata.date", range=m.Range(gte=gte))])
        (f,) = col.resolve_dnf(flt)
        got = (dates >= f.date_lo) & (dates <= f.date_hi)
        assert np.array_equal(got, local_mode.payload_mask(payloads, flt)) and not got[dates == corpus_io.NO_DATE].any()
    for bad in (m.Filter(must=[m.FieldCondition(key="metadata.date", range=m.Range(lte=2 ** 30))]),
                m.Filter(must_not=[m.FieldCondition(key="metadata.date", range=m.Range(gte=0))])):
        with pytest.raises(NotImplementedError):
            col.resolve_dnf(bad)
    assert len(col.resolve_dnf(None)) == 1                              # no date condition: every row passes


@pytest.mark.gpu
def test_load_from_disk_and_search(tmp_path, built):
    from financial_chatbot_llm_b200 import corpus_io, synth
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.engine import Engine
    from oracle import canon
    for uuid_ids, dtype in ((False, "float32"), (True, "bfloat16")):
        spec, hc, pts = make_points(n=1500, dim=128, uuid_ids=uuid_ids)
        d = str(tmp_path / f"c{int(uuid_ids)}")
        corpus_io.write_corpus(d, pts, dim=spec.dim, dtype=dtype, categories=list(synth.CATEGORIES))
        eng = Engine(device=0, max_batch=1, max_k=64)
        col = corpus_io.load_collection(d, eng, chunk_rows=400)
        cl = B200QdrantClient()
        cl.attach_collection(col)
        cf = col.files
        vec = np.asarray(cf.vectors)
        ids = np.array(cf.columns["ids.u64"])
        t0 = int(np.median(cf.columns["dates.i32"]))
        for u in (0, 3, 5):
            q = synth.query_vector(spec, u)
            flt = m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=cf.users[u])),
                                 m.FieldCondition(key="metadata.date", range=m.Range(gte=t0)),
                                 m.FieldCondition(key="metadata.category", match=m.MatchAny(any=["Rent", "Dining", "Fuel", "Other", "Travel"]))])
            pts_out = cl.query_points("transactions", query=q.tolist(), limit=20, query_filter=flt).points
            cmask = sum(1 << cf.manifest["categories"].index(c) for c in ["Rent", "Dining", "Fuel", "Other", "Travel"])
            oi, os_ = canon.search(vec, q, 20, row_begin=int(cf.offsets[u]), row_end=int(cf.offsets[u + 1]), ids=ids,
                                   dates=np.array(cf.columns["dates.i32"]), categories=np.array(cf.columns["categories.u8"]),
                                   date_lo=t0, category_mask=cmask)
            rows = [int(np.flatnonzero(ids == i)[0]) for i in oi]
            want_ids = [cf.point_ids[r] for r in rows] if uuid_ids else [int(i) for i in oi]
            assert [p.id for p in pts_out] == want_ids
            assert [p.score for p in pts_out] == os_.tolist()
            assert all(p.payload == cf.payload(r) for p, r in zip(pts_out, rows))
        eng.close()
