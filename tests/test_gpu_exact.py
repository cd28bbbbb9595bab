"""Exactness by construction: a query whose float32 first pass cannot be CERTIFIED is answered by the all-float64 kernel
(csrc/fcb_exact.cu), and the answer — ids, order, float64 scores, counts — equals the CPU oracle (oracle/canon.c) bit for
bit, the same bar as every other search test. The corpora here are built to defeat the first pass: hundreds of rows
within a few float32 ulps of each other (far more than the 32-candidate margin), exact duplicates whose order only the
point id decides, and rows whose norms overflow float32. An ordinary corpus must NOT trigger the exact path."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def adversarial_corpus(dim=256, n_users=5, seed=9, bf16=False, permute_ids=True):
    """Users 0, 2, 4: ordinary synthetic rows. User 1: 700 rows = one base vector with a few components moved by single
    float32 ulps (cosines differ by ~1e-8..1e-6: far below the first pass's error bound). User 3: 300 rows drawn from only
    6 distinct vectors (exact duplicates: equal scores, order decided by id)."""
    from financial_chatbot_llm_b200 import synth
    spec = synth.SynthSpec(n_rows=6000, dim=dim, n_users=n_users, seed=seed, zipf_a=0.0)
    hc = synth.make_host_corpus(spec)
    rng = np.random.default_rng(seed)
    lo1, hi1 = int(hc.offsets[1]), int(hc.offsets[1]) + 700
    base = hc.vectors[lo1].copy()
    for r in range(lo1, hi1):
        v = base.copy()
        idx = rng.integers(0, dim, size=3)
        steps = rng.integers(-3, 4, size=3)
        v[idx] = (v[idx].view(np.int32) + (steps * (256 * 256 if bf16 else 1)).astype(np.int32)).view(np.float32)
        hc.vectors[r] = v
    lo3, hi3 = int(hc.offsets[3]), int(hc.offsets[3]) + 300
    protos = hc.vectors[lo3:lo3 + 6].copy()
    hc.vectors[lo3:hi3] = protos[rng.integers(0, 6, size=hi3 - lo3)]
    if bf16:
        hc.bf16_bits = synth.f32_to_bf16_bits(hc.vectors).reshape(hc.vectors.shape)
        hc.vectors = synth.bf16_bits_to_f32(hc.bf16_bits).reshape(hc.vectors.shape)
    if permute_ids:
        hc.ids = rng.permutation(spec.n_rows).astype(np.uint64) * 5 + 17
    return spec, hc, base


def oracle(hc, q, k, u, **kw):
    from oracle import canon
    return canon.search(hc.vectors, q, k, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]), ids=hc.ids, **kw)


@pytest.mark.parametrize("bf16", [False, True])
def test_uncertifiable_queries_are_answered_exactly(built, bf16):
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    spec, hc, base = adversarial_corpus(bf16=bf16)
    eng = Engine(device=0, max_batch=8, max_k=128)         # 1200-row segments take the select path for 16 < k <= 128
    loader.collection_from_host(hc, eng, bf16=bf16)
    assert eng.exact_reruns == 0
    seg = lambda u: (int(hc.offsets[u]), int(hc.offsets[u + 1]))
    q_near = base.astype(np.float64) + 1e-3 * synth.query_vector(spec, 1)      # close to user 1's cluster of near-copies
    q_any = synth.query_vector(spec, 2)
    reruns = 0
    for k in (1, 10, 16, 17, 100, 128):                                        # fused scan (<= 16) and the select path
        for u, q in ((1, q_near), (3, q_any), (3, hc.vectors[seg(3)[0] + 7].astype(np.float64))):
            ids, sc, cnt = eng.search_host(q[None, :], [make_filter(*seg(u))], k)
            oi, os_ = oracle(hc, q, k, u)
            assert cnt[0] == len(oi) and ids[0, :cnt[0]].tolist() == oi.tolist(), (k, u)
            assert sc[0, :cnt[0]].tolist() == os_.tolist(), (k, u)
        now = eng.exact_reruns
        if not bf16 or k <= 16:          # (a bf16 ulp is large enough for the select path's 32+ margin to separate the near-copies)
            assert now > reruns, f"k={k}: near-copies / duplicates must have defeated the certificate"
        reruns = now
    # with a date predicate as well (the exact kernel applies the same fused predicate)
    t0 = int(np.median(hc.dates))
    ids, sc, cnt = eng.search_host(q_near[None, :], [make_filter(*seg(1), t0, None)], 10)
    oi, os_ = oracle(hc, q_near, 10, 1, dates=hc.dates, date_lo=t0)
    assert ids[0, :cnt[0]].tolist() == oi.tolist() and sc[0, :cnt[0]].tolist() == os_.tolist()
    eng.close()


def test_batches_mix_certified_and_uncertified_queries_device_api(built):
    """fcb_search (device pointers, asynchronous): the fallback is enqueued behind the scan and re-answers only the
    flagged queries of the batch; the others keep the fast path's results. Inline (batch <= 4) and device-staged filters."""
    import torch
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    spec, hc, base = adversarial_corpus(seed=12)
    eng = Engine(device=0, max_batch=8, max_k=16)
    loader.collection_from_host(hc, eng)
    dev = torch.device("cuda", 0)
    q_near = base.astype(np.float64) + 1e-3 * synth.query_vector(spec, 5)
    for users in ([0, 1, 2], [4, 3, 1, 0, 2, 1, 3]):
        B, k = len(users), 10
        qs = np.stack([q_near if u == 1 else synth.query_vector(spec, 20 + i) for i, u in enumerate(users)])
        fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1])) for u in users]
        ids = torch.empty((B, k), dtype=torch.int64, device=dev)
        sc = torch.empty((B, k), dtype=torch.float64, device=dev)
        cnt = torch.empty(B, dtype=torch.int32, device=dev)
        before = eng.exact_reruns
        eng.search_device(torch.from_numpy(qs).to(dev), fl, k, ids, sc, cnt)
        torch.cuda.synchronize()
        flagged = eng.exact_reruns - before
        assert 1 <= flagged <= sum(u in (1, 3) for u in users), "only the adversarial tenants may need the exact path"
        for i, u in enumerate(users):
            oi, os_ = oracle(hc, qs[i], k, u)
            assert ids[i].cpu().numpy().view(np.uint64).tolist() == oi.tolist(), (users, i)
            assert sc[i].cpu().numpy().tolist() == os_.tolist(), (users, i)
    eng.close()


def test_ordinary_corpus_is_certified_by_the_first_pass(built):
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    spec = synth.SynthSpec(n_rows=40000, dim=1536, n_users=8, seed=3)
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=4, max_k=128)
    loader.collection_from_host(hc, eng)
    for i, (u, k) in enumerate(((0, 10), (1, 16), (0, 100), (5, 10))):
        q = synth.query_vector(spec, 40 + i)
        ids, sc, cnt = eng.search_host(q[None, :], [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]))], k)
        oi, os_ = oracle(hc, q, k, u)
        assert ids[0, :cnt[0]].tolist() == oi.tolist() and sc[0, :cnt[0]].tolist() == os_.tolist()
    assert eng.exact_reruns == 0, "continuous embeddings must not need the exact path"
    eng.close()


def test_rows_outside_the_float32_range_send_every_query_to_the_exact_path(built):
    """A row scaled by 1e25 overflows the float32 sum of squares: the first pass has no error bound on such a shard, so
    fcb_load_shard notices and every query is answered by the float64 kernel — still equal to the oracle."""
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    spec = synth.SynthSpec(n_rows=3000, dim=128, n_users=3, seed=8)
    hc = synth.make_host_corpus(spec)
    hc.vectors[int(hc.offsets[1]) + 5] *= np.float32(1e25)
    hc.vectors[int(hc.offsets[1]) + 9] *= np.float32(1e-28)
    eng = Engine(device=0, max_batch=2, max_k=64)
    loader.collection_from_host(hc, eng)
    for k in (5, 40):
        for u in (0, 1):
            q = synth.query_vector(spec, 60 + u)
            before = eng.exact_reruns
            ids, sc, cnt = eng.search_host(q[None, :], [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]))], k)
            oi, os_ = oracle(hc, q, k, u)
            assert ids[0, :cnt[0]].tolist() == oi.tolist() and sc[0, :cnt[0]].tolist() == os_.tolist(), (k, u)
            assert eng.exact_reruns == before + 1
    eng.close()
