"""BASELINE.json configs[0] exactly as SURVEY.md §8(d) states it: 100 000 x 1536 float32, cosine, 1 000 Zipf users,
top-10, `user_id` filter, 1 000 single queries (batch 1) — the configuration the reference itself runs (qdrant-client local
mode through tools/qdrant_tool.py:105-112,145-153). Every query's ids, order, count and float64 scores must equal the C
oracle bit for bit; a sample of them is also checked against the numpy local-mode model (which scores ALL 100k rows per
query, as local mode does) at north_star's 1e-5 relative tolerance for float32 storage."""
import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]
REL_TOL_F32 = 1e-5


def test_config0_1000_single_queries_match_the_oracle(built):
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon, local_mode
    spec = synth.SynthSpec(n_rows=100_000, dim=1536, n_users=1000, seed=20240607)     # Zipf segment sizes (zipf_a = 1.1)
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=1)                                               # INTEGRATION.md's recipe
    col = loader.collection_from_host(hc, eng)
    client = B200QdrantClient(url="unused")
    client.attach_collection(col)
    rng = np.random.default_rng(0)
    # 1 000 queries: the 40 largest tenants, then random tenants weighted by size (whoever owns more rows asks more)
    sizes = np.diff(hc.offsets)
    users = np.concatenate([np.argsort(-sizes)[:40], rng.choice(spec.n_users, size=960, p=sizes / sizes.sum())])
    K = 10
    n_local = 0
    for i, u in enumerate(users):
        u = int(u)
        q = synth.query_vector(spec, i)
        rb, re_ = int(hc.offsets[u]), int(hc.offsets[u + 1])
        oi, os_ = canon.search(hc.vectors, q, K, row_begin=rb, row_end=re_, ids=hc.ids)
        if i % 2 == 0:                                                                # through the C-ABI host call
            ids, sc, cnt = eng.search_host(q[None, :], [make_filter(rb, re_)], K)
            got_ids, got_sc = ids[0, :cnt[0]].tolist(), sc[0, :cnt[0]].tolist()
        else:                                                                         # through the qdrant-shaped adapter
            pts = client.query_points("transactions", query=q.tolist(), limit=K, query_filter=m.Filter(must=[
                m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[u]))])).points
            got_ids, got_sc = [p.id for p in pts], [p.score for p in pts]
            assert all(p.payload["metadata"]["user_id"] == hc.users[u] for p in pts)
        assert got_ids == oi.tolist(), (i, u)
        assert got_sc == os_.tolist(), (i, u)
        assert len(got_ids) == min(K, re_ - rb)
        if i % 50 == 0:                                                               # local-mode model: all rows scored, mask applied
            mask = local_mode.column_mask(spec.n_rows, hc.offsets, u, hc.dates, None, None, None, None)
            li, ls = local_mode.search_columns(hc.vectors.copy(), hc.ids, mask, q.tolist(), K, inplace_renorm=True)
            assert [int(x) for x in li] == got_ids, (i, u)
            assert np.allclose(np.asarray(ls, dtype=np.float64), np.asarray(got_sc), rtol=REL_TOL_F32, atol=0)
            n_local += 1
    assert n_local == 20
    assert eng.exact_reruns == 0, "continuous embeddings: every query certified by the first pass"
    eng.close()
