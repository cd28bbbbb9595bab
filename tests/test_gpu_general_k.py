"""General-k path (16 < k <= 10000; the tool's default limit is 10000, tools/qdrant_tool.py:145):
radix-select of float32 keys + canonical float64 rescore, against the C oracle, bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(built):
    import torch
    assert torch.cuda.is_available()
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine
    spec = synth.SynthSpec(n_rows=40000, dim=512, n_users=9, seed=808, whale_rows=25000)
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=4, max_k=10000)
    col = loader.collection_from_host(hc, eng)
    yield hc, eng, col
    eng.close()


def check(eng, hc, qi, k, user, **kw):
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import make_filter
    from oracle import canon
    q = synth.query_vector(hc.spec, qi)
    rb, re_ = int(hc.offsets[user]), int(hc.offsets[user + 1])
    ids, sc, cnt = eng.search_host(q[None, :], [make_filter(rb, re_, kw.get("date_lo"), kw.get("date_hi"), kw.get("category_mask"))], k)
    oi, os_ = canon.search(hc.vectors, q, k, row_begin=rb, row_end=re_, ids=hc.ids, dates=hc.dates,
                           categories=hc.categories, **kw)
    n = int(cnt[0])
    assert n == len(oi)
    assert ids[0, :n].tolist() == oi.tolist()
    assert sc[0, :n].tolist() == os_.tolist()
    assert (ids[0, n:] == np.iinfo(np.uint64).max).all() and np.isneginf(sc[0, n:]).all()
    return n


def test_k_sweep_on_whale_segment(env):
    hc, eng, _ = env
    for k in (17, 64, 100, 1000, 10000):
        assert check(eng, hc, k, k, 0) == k
    assert eng.last_search_launches == 15               # key scan, 8 radix passes, compact, re-score, chunk sort, rank merge,
                                                        # certificate, flag-gated float64 path


def test_k_larger_than_segment_and_filters(env):
    hc, eng, _ = env
    t0 = int(np.percentile(hc.dates, 70))
    n = check(eng, hc, 1, 10000, 3)                         # default limit on a small tenant
    assert n == int(hc.offsets[4] - hc.offsets[3])
    check(eng, hc, 2, 500, 0, date_lo=t0)
    check(eng, hc, 3, 2000, 0, date_lo=t0, category_mask=(1 << 4) | (1 << 5))
    check(eng, hc, 4, 50, 5, category_mask=1 << 0)
    assert check(eng, hc, 5, 100, 2, date_lo=2 ** 40) == 0  # nothing passes


def test_segments_that_fit_the_candidate_buffer_take_the_direct_path(env):
    """The tool's default call (limit 10000) against an ordinary tenant: every passing row is scored in float64
    straight away (no key scan, no radix select), then sorted — ONE launch up to 1024 rows, 3 above."""
    hc, eng, _ = env
    sizes = np.diff(hc.offsets)
    small = int(np.argmin(np.where(sizes > 0, sizes, 1 << 30)))
    big = int(np.argmax(np.where(sizes <= 10000, sizes, -1)))
    assert sizes[big] > 1024
    t0 = int(np.percentile(hc.dates, 40))
    for k in (10000, 1500, 100, 17):
        assert check(eng, hc, 7, k, big) == min(k, sizes[big])
        assert eng.last_search_launches == 3
        check(eng, hc, 8, k, big, date_lo=t0, category_mask=0b1011011)
    if sizes[small] <= 1024:
        for rep in range(3):                                  # back to back: the state is left clean by the launch itself
            assert check(eng, hc, 9 + rep, 10000, small) == sizes[small]
            assert eng.last_search_launches == 1
        check(eng, hc, 9, 20, small, date_lo=t0)
        check(eng, hc, 7, 64, big)                              # and the other paths still start from a clean state
        assert check(eng, hc, 9, 10000, small) == sizes[small]


def test_batch_of_general_queries(env):
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import make_filter
    from oracle import canon
    hc, eng, _ = env
    qs = np.stack([synth.query_vector(hc.spec, 50 + i) for i in range(3)])
    fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1])) for u in (0, 4, 8)]
    ids, sc, cnt = eng.search_host(qs, fl, 300)
    for i, u in enumerate((0, 4, 8)):
        oi, os_ = canon.search(hc.vectors, qs[i], 300, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
        assert ids[i, :cnt[i]].tolist() == oi.tolist() and sc[i, :cnt[i]].tolist() == os_.tolist()


def test_ties_across_the_cut_with_permuted_ids(built):
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    spec = synth.SynthSpec(n_rows=6000, dim=128, n_users=1, seed=6)
    vec = synth.vectors_f32(spec, 0, 6000)
    q = synth.query_vector(spec, 1)
    order = np.argsort(-canon.scores(vec, q))
    vec[np.arange(7, 6000, 13)] = vec[order[40]]            # ~460 copies of the 41st-best row
    rng = np.random.default_rng(3)
    ids = rng.permutation(6000).astype(np.uint64) + np.uint64(10 ** 12)
    rank = np.empty(6000, dtype=np.int64)
    rank[np.argsort(ids)] = np.arange(6000)
    eng = Engine(device=0, max_batch=1, max_k=1000)
    dev = torch.device("cuda")
    eng.load_shard(torch.from_numpy(vec).to(dev), ids=torch.from_numpy(ids.view(np.int64)).to(dev),
                   id_rank=torch.from_numpy(rank.astype(np.int32)).to(dev))
    for k in (45, 100, 400, 1000):
        gi, gs, cnt = eng.search_host(q[None, :], [make_filter(0, 6000)], k)
        oi, os_ = canon.search(vec, q, k, ids=ids)
        assert cnt[0] == k and gi[0].tolist() == oi.tolist() and gs[0].tolist() == os_.tolist()
    eng.close()
    # the same corpus through the direct path (the segment fits the candidate buffer): ties inside and across sorted chunks
    eng = Engine(device=0, max_batch=1, max_k=10000)
    eng.load_shard(torch.from_numpy(vec).to(dev), ids=torch.from_numpy(ids.view(np.int64)).to(dev),
                   id_rank=torch.from_numpy(rank.astype(np.int32)).to(dev))
    for k in (45, 1000, 6000, 10000):
        gi, gs, cnt = eng.search_host(q[None, :], [make_filter(0, 6000)], k)
        oi, os_ = canon.search(vec, q, k, ids=ids)
        n = min(k, 6000)
        assert cnt[0] == n and gi[0, :n].tolist() == oi.tolist() and gs[0, :n].tolist() == os_.tolist()
        assert eng.last_search_launches == 3
    eng.close()


def test_bf16_general_k(built):
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    spec = synth.SynthSpec(n_rows=12000, dim=1024, n_users=3, seed=12)
    hc = synth.make_host_corpus(spec, bf16=True)
    eng = Engine(device=0, max_batch=1, max_k=2048)
    loader.collection_from_host(hc, eng, bf16=True)
    q = synth.query_vector(spec, 0)
    ids, sc, cnt = eng.search_host(q[None, :], [make_filter(0, int(hc.offsets[1]))], 2048)
    oi, os_ = canon.search(hc.bf16_bits, q, 2048, row_begin=0, row_end=int(hc.offsets[1]))
    assert ids[0, :cnt[0]].tolist() == oi.tolist() and sc[0, :cnt[0]].tolist() == os_.tolist()
    eng.close()


def test_batch_of_small_tenants_is_one_launch(env):
    """A micro-batch of default-limit calls against ordinary tenants (<= 1024 rows each): one launch scores, sorts and
    writes every query (blockIdx.y = query); repeated calls prove that the per-query counters are left clean."""
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import make_filter
    from oracle import canon
    hc, eng, _ = env
    users = [5, 8, 6, 7]
    assert all(hc.offsets[u + 1] - hc.offsets[u] <= 1024 for u in users)
    t0 = int(np.percentile(hc.dates, 55))
    for rep, k in enumerate((10000, 50, 10000)):
        qs = np.stack([synth.query_vector(hc.spec, 70 + rep * 4 + i) for i in range(4)])
        kw = [dict(), dict(date_lo=t0), dict(category_mask=0b110101), dict(date_lo=2 ** 40)]     # the last one: nothing passes
        fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]), w.get("date_lo"), None, w.get("category_mask"))
              for u, w in zip(users, kw)]
        ids, sc, cnt = eng.search_host(qs, fl, k)
        assert eng.last_search_launches == 1
        for i, (u, w) in enumerate(zip(users, kw)):
            oi, os_ = canon.search(hc.vectors, qs[i], k, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]),
                                   ids=hc.ids, dates=hc.dates, categories=hc.categories, **w)
            n = int(cnt[i])
            assert n == len(oi) and ids[i, :n].tolist() == oi.tolist() and sc[i, :n].tolist() == os_.tolist()
            assert (ids[i, n:] == np.iinfo(np.uint64).max).all() and np.isneginf(sc[i, n:]).all()
    assert cnt[3] == 0
