"""GPU parity tests: the CUDA path, called through the C-ABI, against the oracle on the same
seeded inputs. Integer/index results (ids, counts, order) are bit-exact; float64 scores are
compared bit-exact against the C restatement (same pinned order of operations) and within
1e-5 relative (north_star's fp32-storage tolerance; 2e-2 for bf16 storage) against the numpy
local-mode model."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_TOL_F32 = 1e-5       # north_star: scores within 1e-5 relative for fp32 storage
REL_TOL_BF16 = 2e-2      # north_star: 2e-2 for bf16 storage (rescored from the stored values)


@pytest.fixture(scope="module")
def env(built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine
    spec = synth.SynthSpec(n_rows=30000, dim=1536, n_users=25, seed=77)
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=16, max_k=16)
    col = loader.collection_from_host(hc, eng)
    yield hc, eng, col
    eng.close()


def oracle_search(hc, q, k, user=None, **kw):
    from oracle import canon
    rb, re_ = (0, hc.spec.n_rows) if user is None else (int(hc.offsets[user]), int(hc.offsets[user + 1]))
    return canon.search(hc.vectors, q, k, row_begin=rb, row_end=re_, ids=hc.ids, dates=hc.dates,
                        categories=hc.categories, **kw)


def run(eng, hc, q, k, user, date_lo=None, date_hi=None, category_mask=None):
    from financial_chatbot_llm_b200.engine import make_filter
    rb, re_ = (0, hc.spec.n_rows) if user is None else (int(hc.offsets[user]), int(hc.offsets[user + 1]))
    ids, sc, cnt = eng.search_host(q[None, :], [make_filter(rb, re_, date_lo, date_hi, category_mask)], k)
    n = int(cnt[0])
    return ids[0, :n], sc[0, :n], n, ids[0], sc[0]


def test_device_generator_is_bit_identical_to_host(built):
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import synth_columns, synth_vectors
    spec = synth.SynthSpec(n_rows=5000, dim=1536, n_users=3, seed=123)
    for dtype, host in ((torch.float32, synth.vectors_f32(spec, 1000, 1300)),
                        (torch.bfloat16, synth.vectors_bf16_bits(spec, 1000, 1300))):
        out = torch.empty((300, spec.dim), dtype=dtype, device="cuda")
        synth_vectors(out, 1000, spec.seed, spec.n_clusters)
        torch.cuda.synchronize()
        got = out.cpu().numpy() if dtype == torch.float32 else out.view(torch.int16).cpu().numpy().view(np.uint16)
        assert np.array_equal(got, host.reshape(300, spec.dim))
    d = torch.empty(4000, dtype=torch.int32, device="cuda")
    c = torch.empty(4000, dtype=torch.uint8, device="cuda")
    v = torch.empty(4000, dtype=torch.int32, device="cuda")
    synth_columns(500, 4000, spec.seed, spec.n_categories, spec.date_span_seconds, dates=d, categories=c, cents=v)
    torch.cuda.synchronize()
    assert np.array_equal(d.cpu().numpy(), synth.dates(spec, 500, 4500))
    assert np.array_equal(c.cpu().numpy(), synth.categories(spec, 500, 4500))
    assert np.array_equal(v.cpu().numpy(), synth.cents(spec, 500, 4500))


def test_user_filter_top10_bit_exact(env):
    from financial_chatbot_llm_b200 import synth
    hc, eng, _ = env
    for qi, user in [(0, 0), (1, 3), (2, 24), (3, 11), (4, None)]:
        q = synth.query_vector(hc.spec, qi)
        ids, sc, n, _, _ = run(eng, hc, q, 10, user)
        oi, os_ = oracle_search(hc, q, 10, user)
        assert n == len(oi) == 10
        assert ids.tolist() == oi.tolist()
        assert sc.tolist() == os_.tolist()          # float64, same order of operations => identical bits
        assert eng.last_search_launches == 1


def test_date_and_category_predicates_fused_in_scan(env):
    from financial_chatbot_llm_b200 import synth
    hc, eng, _ = env
    t0, t1 = int(np.percentile(hc.dates, 40)), int(np.percentile(hc.dates, 90))
    cmask = (1 << 2) | (1 << 7) | (1 << 15)
    q = synth.query_vector(hc.spec, 5)
    for user, kw in [(0, dict(date_lo=t0)), (1, dict(date_lo=t0, date_hi=t1)), (2, dict(category_mask=cmask)),
                     (0, dict(date_lo=t0, category_mask=cmask)), (5, dict(date_hi=t0 - 1)),
                     (None, dict(date_lo=t1, category_mask=1 << 3))]:
        ids, sc, n, _, _ = run(eng, hc, q, 10, user, **kw)
        oi, os_ = oracle_search(hc, q, 10, user, **kw)
        assert ids.tolist() == oi.tolist() and sc.tolist() == os_.tolist(), (user, kw)


def test_edge_cases_empty_ragged_and_k_bounds(env):
    from financial_chatbot_llm_b200 import _lib, synth
    from financial_chatbot_llm_b200.engine import make_filter
    hc, eng, _ = env
    q = synth.query_vector(hc.spec, 6)
    # unknown user / empty segment
    ids, sc, cnt = eng.search_host(q[None, :], [make_filter(0, 0)], 10)
    assert cnt[0] == 0 and (ids[0] == np.iinfo(np.uint64).max).all() and np.isneginf(sc[0]).all()
    # predicate nothing passes
    _, _, n, _, _ = run(eng, hc, q, 10, 4, date_lo=2**40)
    assert n == 0
    # segment smaller than k: counts < k, tail padded
    small = int(np.argmin(np.diff(hc.offsets)))
    size = int(hc.offsets[small + 1] - hc.offsets[small])
    tiny_hi = int(hc.offsets[small]) + 3
    ids, sc, cnt = eng.search_host(q[None, :], [make_filter(int(hc.offsets[small]), tiny_hi)], 16)
    assert cnt[0] == 3 and (ids[0, 3:] == np.iinfo(np.uint64).max).all()
    from oracle import canon
    oi, os_ = canon.search(hc.vectors, q, 16, row_begin=int(hc.offsets[small]), row_end=tiny_hi)
    assert ids[0, :3].tolist() == oi.tolist() and sc[0, :3].tolist() == os_.tolist()
    assert size >= 3
    for k in (1, 2, 16):
        ids, sc, n, _, _ = run(eng, hc, q, k, 7)
        oi, os_ = oracle_search(hc, q, k, 7)
        assert ids.tolist() == oi.tolist() and sc.tolist() == os_.tolist()
    with pytest.raises(_lib.FcbError):
        eng.search_host(q[None, :], [make_filter(0, 10)], 17)          # k > max_k
    with pytest.raises(_lib.FcbError):
        eng.search_host(np.zeros((17, hc.spec.dim)), [make_filter(0, 10)] * 17, 5)   # batch > max_batch


def test_batched_queries_inline_and_staged_filters(env):
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import make_filter
    hc, eng, _ = env
    t0 = int(np.percentile(hc.dates, 50))
    for B in (3, 4, 5, 16):
        qs = np.stack([synth.query_vector(hc.spec, 100 + i) for i in range(B)])
        users = [(7 * i + 1) % 25 for i in range(B)]
        fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]), t0 if i % 2 else None) for i, u in enumerate(users)]
        ids, sc, cnt = eng.search_host(qs, fl, 10)
        for i, u in enumerate(users):
            oi, os_ = oracle_search(hc, qs[i], 10, u, **({"date_lo": t0} if i % 2 else {}))
            assert ids[i, :cnt[i]].tolist() == oi.tolist() and sc[i, :cnt[i]].tolist() == os_.tolist()


def test_device_pointer_entry_point_async(env):
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import make_filter
    hc, eng, _ = env
    qs = np.stack([synth.query_vector(hc.spec, 200 + i) for i in range(2)])
    dq = torch.from_numpy(qs).cuda()
    out_ids = torch.empty((2, 10), dtype=torch.int64, device="cuda")
    out_sc = torch.empty((2, 10), dtype=torch.float64, device="cuda")
    out_cnt = torch.empty(2, dtype=torch.int32, device="cuda")
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        eng.search_device(dq, [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1])) for u in (2, 9)], 10,
                          out_ids, out_sc, out_cnt)
    st.synchronize()
    for i, u in enumerate((2, 9)):
        oi, os_ = oracle_search(hc, qs[i], 10, u)
        assert out_ids[i].cpu().numpy().view(np.uint64).tolist() == oi.tolist()
        assert out_sc[i].cpu().numpy().tolist() == os_.tolist()


def test_ties_resolved_by_ascending_id_with_permuted_ids(built):
    """Exact duplicates (equal scores) straddling the k-th place; ids are NOT monotone in row
    order, so the id_rank column carries the tie order into the first-pass selection."""
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    spec = synth.SynthSpec(n_rows=4000, dim=256, n_users=2, seed=5)
    vec = synth.vectors_f32(spec, 0, 4000)
    q = synth.query_vector(spec, 3)
    best = int(np.argmax(canon.scores(vec, q)))
    dup_rows = np.arange(100, 4000, 61)                      # 64 copies of the best row, spread out
    vec[dup_rows] = vec[best]
    rng = np.random.default_rng(1)
    ids = rng.permutation(4000).astype(np.uint64) * np.uint64(1000003) + np.uint64(17)
    rank = np.empty(4000, dtype=np.int64)
    rank[np.argsort(ids)] = np.arange(4000)
    eng = Engine(device=0, max_batch=1, max_k=16)
    dev = torch.device("cuda")
    eng.load_shard(torch.from_numpy(vec).to(dev), ids=torch.from_numpy(ids.view(np.int64)).to(dev),
                   id_rank=torch.from_numpy(rank.astype(np.int32)).to(dev))
    for k in (1, 10, 16):
        got_ids, got_sc, cnt = eng.search_host(q[None, :], [make_filter(0, 4000)], k)
        oi, os_ = canon.search(vec, q, k, ids=ids)
        assert cnt[0] == k and got_ids[0].tolist() == oi.tolist() and got_sc[0].tolist() == os_.tolist()
        assert len(set(got_sc[0].tolist())) == 1                 # all ties
    eng.close()


def test_bf16_storage_with_float64_rescore(built):
    import torch
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon, local_mode
    spec = synth.SynthSpec(n_rows=20000, dim=1024, n_users=12, seed=31)
    hc = synth.make_host_corpus(spec, bf16=True)
    eng = Engine(device=0, max_batch=4, max_k=16)
    loader.collection_from_host(hc, eng, bf16=True)
    t0 = int(np.percentile(hc.dates, 30))
    for qi, u in [(0, 0), (1, 5), (2, 11)]:
        q = synth.query_vector(spec, qi)
        ids, sc, cnt = eng.search_host(q[None, :], [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]), t0)], 10)
        oi, os_ = canon.search(hc.bf16_bits, q, 10, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]),
                               dates=hc.dates, date_lo=t0)
        assert ids[0].tolist() == oi.tolist() and sc[0].tolist() == os_.tolist()
        # against the numpy local-mode model over the bf16-rounded values (the corpus IS those values)
        mask = local_mode.column_mask(spec.n_rows, hc.offsets, u, hc.dates, t0, None, None, None)
        i_np, s_np = local_mode.search_columns(hc.vectors, hc.ids, mask, q, 10)
        assert ids[0].tolist() == i_np.tolist()
        assert np.all(np.abs(sc[0] - s_np) <= REL_TOL_BF16 * np.abs(s_np))
    eng.close()


def test_adapter_query_points_against_local_mode_model(env):
    """The qdrant-shaped call through the adapter vs the numpy model of local mode."""
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from oracle import local_mode
    hc, eng, col = env
    cl = B200QdrantClient(url="unused", api_key="unused")
    cl.attach_collection(col)
    t0 = int(np.percentile(hc.dates, 60))
    for qi, u in [(10, 1), (11, 8), (12, 20)]:
        q = synth.query_vector(hc.spec, qi)
        flt = m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[u])),
                             m.FieldCondition(key="metadata.date", range=m.Range(gte=t0)),
                             m.FieldCondition(key="metadata.category", match=m.MatchAny(any=["Rent", "Fuel", "Dining", "Other"]))])
        pts = cl.query_points(collection_name="transactions", query=q.tolist(), limit=10,
                              search_params=m.SearchParams(hnsw_ef=128, exact=False), query_filter=flt).points
        cmask = sum(1 << synth.CATEGORIES.index(c) for c in ["Rent", "Fuel", "Dining", "Other"])
        mask = local_mode.column_mask(hc.spec.n_rows, hc.offsets, u, hc.dates, t0, None, hc.categories, cmask)
        i_np, s_np = local_mode.search_columns(hc.vectors.copy(), hc.ids, mask, q, 10, inplace_renorm=True)
        assert [p.id for p in pts] == i_np.tolist()
        got = np.array([p.score for p in pts])
        assert np.all(np.abs(got - s_np) <= REL_TOL_F32 * np.abs(s_np))
        for p in pts:
            md = p.payload["metadata"]
            assert md["user_id"] == hc.users[u] and md["date"] >= t0 and md["category"] in ("Rent", "Fuel", "Dining", "Other")
            assert p.payload == hc.payload(p.id) and p.version == 0 and p.vector is None


def test_retrieve_transactions_end_to_end(env):
    import datetime as dt
    from financial_chatbot_llm_b200 import synth, tools
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from oracle import canon
    hc, eng, col = env
    cl = B200QdrantClient()
    cl.attach_collection(col)
    q = synth.query_vector(hc.spec, 42)
    now = dt.datetime.fromtimestamp(synth.DATE_BASE + 700 * 86400)
    tools.configure(cl, lambda text: q.tolist(), now=lambda: now)
    out = tools.retrieve_transactions(user_id=hc.users[6], num_transactions=10, time_period_days=400)
    t0 = int((now - dt.timedelta(days=400)).timestamp())
    oi, _ = canon.search(hc.vectors, q, 10, row_begin=int(hc.offsets[6]), row_end=int(hc.offsets[7]),
                         dates=hc.dates, date_lo=t0)
    assert out == [hc.payload(int(r))["page_content"] for r in oi] and len(out) == 10
    assert tools.retrieve_transactions(user_id="no-such-user", num_transactions=5) == []


def test_should_and_must_not_filters_against_local_mode_model(env):
    """OR / NOT filters are rewritten into several conjunctions that ride in one grouped launch and are
    merged on the host (client.py); the selected points must be what the filter model selects."""
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from oracle import local_mode
    hc, eng, col = env
    cl = B200QdrantClient()
    cl.attach_collection(col)
    payloads = [hc.payload(r) for r in range(hc.spec.n_rows)]
    t1, t2 = int(np.percentile(hc.dates, 30)), int(np.percentile(hc.dates, 70))

    def U(u): return m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[u]))
    def D(**kw): return m.FieldCondition(key="metadata.date", range=m.Range(**kw))
    def C(*v): return m.FieldCondition(key="metadata.category", match=m.MatchAny(any=list(v)))
    filters = [
        m.Filter(must=[U(3)], must_not=[D(gte=t1, lte=t2)]),                       # before OR after the window
        m.Filter(must_not=[U(0), U(7)]),                                           # everyone else's rows
        m.Filter(should=[U(2), U(9), U(17)], must_not=[C("Rent", "Fuel")]),
        m.Filter(must=[D(gte=t1)], should=[C("Dining"), D(lte=t2)], must_not=[U(5)]),   # overlapping conjuncts
        m.Filter(must=[U(4)], must_not=[U(4)]),                                    # nothing
    ]
    for n_f, flt in enumerate(filters):
        q = synth.query_vector(hc.spec, 50 + n_f)
        pts = cl.query_points(collection_name="transactions", query=q.tolist(), limit=12, query_filter=flt).points
        mask = local_mode.payload_mask(payloads, flt)
        i_np, s_np = local_mode.search_columns(hc.vectors.copy(), hc.ids, mask, q, 12, inplace_renorm=True)
        assert [p.id for p in pts] == i_np.tolist(), n_f
        got = np.array([p.score for p in pts])
        assert np.all(np.abs(got - s_np) <= REL_TOL_F32 * np.abs(s_np))
        assert all(local_mode.check_filter(flt, p.payload) for p in pts)
    assert cl.query_points("transactions", query=q.tolist(), limit=5, query_filter=filters[-1]).points == []
