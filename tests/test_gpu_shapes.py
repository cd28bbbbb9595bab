"""Shapes and layouts off the benchmark path: every (storage, dim) route of the (R,U) table, tiny dims,
shard row_base / id_base offsets, ids without id_rank (ascending), a shard that holds only part of a segment."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dim,bf16", [(8, False), (36, False), (128, False), (384, False), (768, False), (1024, False),
                                      (3072, False), (8, True), (136, True), (512, True), (768, True), (1024, True), (2048, True)])
def test_dims_and_storage_types(built, dim, bf16):
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    spec = synth.SynthSpec(n_rows=2500, dim=dim, n_users=4, seed=dim)
    hc = synth.make_host_corpus(spec, bf16=bf16)
    eng = Engine(device=0, max_batch=2, max_k=40)
    loader.collection_from_host(hc, eng, bf16=bf16)
    vec = hc.bf16_bits if bf16 else hc.vectors
    t0 = int(np.median(hc.dates))
    for k in (10, 40):
        for u, dlo in ((0, None), (3, t0)):
            q = synth.query_vector(spec, u + k)
            rb, re_ = int(hc.offsets[u]), int(hc.offsets[u + 1])
            ids, sc, cnt = eng.search_host(q[None, :], [make_filter(rb, re_, dlo)], k)
            oi, os_ = canon.search(vec, q, k, row_begin=rb, row_end=re_, ids=hc.ids, dates=hc.dates, date_lo=dlo)
            assert ids[0, :cnt[0]].tolist() == oi.tolist() and sc[0, :cnt[0]].tolist() == os_.tolist()
    eng.close()


def test_partial_shard_row_base_and_implicit_ids(built):
    """A shard holding global rows [700, 2300) of a 3000-row collection, ids implicit (= global row)."""
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    spec = synth.SynthSpec(n_rows=3000, dim=64, n_users=5, seed=9)
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=1, max_k=64)
    loader.collection_from_host(hc, eng, with_ids=False, row_lo=700, row_hi=2300)
    for u in range(5):
        q = synth.query_vector(spec, u)
        rb, re_ = int(hc.offsets[u]), int(hc.offsets[u + 1])
        for k in (10, 64):
            ids, sc, cnt = eng.search_host(q[None, :], [make_filter(rb, re_)], k)
            lo, hi = max(rb, 700), min(re_, 2300)
            oi, os_ = (np.zeros(0, np.uint64), np.zeros(0)) if hi <= lo else canon.search(hc.vectors, q, k, row_begin=lo, row_end=hi)
            assert ids[0, :cnt[0]].tolist() == oi.tolist() and sc[0, :cnt[0]].tolist() == os_.tolist()
    eng.close()
