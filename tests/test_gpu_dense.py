"""Batched tensor-core path (tcgen05/TMEM/TMA; BASELINE config 3): many queries sharing one segment of a
bf16 shard. First pass in bf16 on the tensor cores, canonical float64 rescore, certificate + exact
re-run; results must equal the C oracle bit for bit (ids, float64 scores, counts)."""
import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def run_case(spec, whale_rows, batches, k=10, date_frac=None, mutate=None, expect_redo=False):
    import torch
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    hc = synth.make_host_corpus(spec, bf16=True)
    if mutate is not None:
        mutate(hc)
    eng = Engine(device=0, max_batch=256, max_k=16)
    loader.collection_from_host(hc, eng, bf16=True)
    rb, re_ = int(hc.offsets[0]), int(hc.offsets[1])
    assert re_ - rb == whale_rows >= 65536
    dlo = None if date_frac is None else int(np.percentile(hc.dates, 100 * date_frac))
    for B in batches:
        qs = np.stack([synth.query_vector(spec, 1000 + i) for i in range(B)])
        ids, sc, cnt = eng.search_host(qs, [make_filter(rb, re_, dlo, None)] * B, k)
        assert eng.last_search_launches == 4          # pack + tcgen05 pass + finalize + (flag-gated) exact re-run
        redo = eng.last_dense_redo(B)                 # queries the tensor-core pass could not certify
        assert (redo > 0) if expect_redo else (redo == 0), redo
        for b in range(B):
            oi, os_ = canon.search(hc.bf16_bits, qs[b], k, row_begin=rb, row_end=re_, ids=hc.ids, dates=hc.dates, date_lo=dlo)
            assert cnt[b] == len(oi) == k
            assert ids[b].tolist() == oi.tolist(), (B, b)
            assert sc[b].tolist() == os_.tolist(), (B, b)
    eng.close()


def test_dense_small_dim_various_batches(built):
    from financial_chatbot_llm_b200 import synth
    spec = synth.SynthSpec(n_rows=72000, dim=128, n_users=5, seed=51, whale_rows=70001)
    run_case(spec, 70001, (32, 40, 256))


def test_dense_dim1536_with_shared_date_filter(built):
    from financial_chatbot_llm_b200 import synth
    spec = synth.SynthSpec(n_rows=66600, dim=1536, n_users=3, seed=52, whale_rows=66000)
    run_case(spec, 66000, (64,), k=10, date_frac=0.3)


def test_dense_uncertified_queries_fall_back_to_exact_scan(built):
    """~200 rows within 1e-4 of the best score: the bf16-query first pass cannot certify the top-10,
    so the flag-gated exact scan must produce the answer (and ties by id must hold)."""
    from financial_chatbot_llm_b200 import synth
    from oracle import canon
    spec = synth.SynthSpec(n_rows=70000, dim=256, n_users=2, seed=53, whale_rows=69000)

    def mutate(hc):
        q = synth.query_vector(spec, 1000)
        best = int(np.argmax(canon.scores(hc.bf16_bits[:69000], q)))
        rows = np.arange(11, 69000, 345)
        hc.bf16_bits[rows] = hc.bf16_bits[best]
        hc.bf16_bits[rows[::2], 5] ^= 1                      # half of them differ in one low bit
        hc.vectors[:] = synth.bf16_bits_to_f32(hc.bf16_bits).reshape(hc.vectors.shape)
    run_case(spec, 69000, (32,), k=10, mutate=mutate, expect_redo=True)


def test_mixed_filters_use_the_grouped_scan(built):
    """Different filters per query -> not a dense contraction -> grouped GEMV launch (1 kernel)."""
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    spec = synth.SynthSpec(n_rows=70000, dim=128, n_users=40, seed=54)
    hc = synth.make_host_corpus(spec, bf16=True)
    eng = Engine(device=0, max_batch=64, max_k=16)
    loader.collection_from_host(hc, eng, bf16=True)
    B = 48
    qs = np.stack([synth.query_vector(spec, i) for i in range(B)])
    users = [i % 40 for i in range(B)]
    ids, sc, cnt = eng.search_host(qs, [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1])) for u in users], 10)
    assert eng.last_search_launches == 1
    for b, u in enumerate(users):
        oi, os_ = canon.search(hc.bf16_bits, qs[b], 10, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
        assert ids[b, :cnt[b]].tolist() == oi.tolist() and sc[b, :cnt[b]].tolist() == os_.tolist()
    eng.close()
