"""fcb_search_noalloc — the entry point for callers that cannot allocate GPU memory (a worker on
a GPU whose HBM a resident corpus owner has filled; a request thread that must not hit
cudaMalloc's device-wide synchronisation). It keeps its scratch in a fixed array inside the
library, so this test checks, through the C-ABI, that it returns exactly what the allocating
route (fcb_open + fcb_load_shard + fcb_search) returns, and what the CPU oracle (oracle/canon.c)
returns — ids, float64 scores and counts equal bit for bit, the tolerance every search parity
test here uses — on two calls IN A ROW, because the second call starts from whatever the first
left in the shared scratch.

Named test_zz_* so pytest collects it last: a run stopped at the first failure has then already
run every other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_matches_engine_on_two_consecutive_calls(built):
    import torch
    assert torch.cuda.is_available()
    from financial_chatbot_llm_b200 import _lib, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter, search_noalloc
    from oracle import canon

    dev = torch.device("cuda", 0)
    spec = synth.SynthSpec(n_rows=24000, dim=1536, n_users=6, seed=404)
    hc = synth.make_host_corpus(spec)
    vec = torch.from_numpy(hc.vectors).to(dev)
    dates = torch.from_numpy(hc.dates).to(dev)
    cats = torch.from_numpy(hc.categories).to(dev)

    eng = Engine(device=0, max_batch=1, max_k=16)
    eng.load_shard(vec, dates=dates, categories=cats)
    shard = _lib.FcbShard(vec.data_ptr(), None, None, dates.data_ptr(), cats.data_ptr(), None, None,
                          spec.n_rows, 0, 0, spec.dim, _lib.FCB_DTYPE_F32)

    t0 = int(np.percentile(hc.dates, 35))
    calls = [  # (query index, user, k, date_lo, category mask): two calls back to back, then a ragged one
        (0, 0, 10, t0, None),
        (1, 3, 16, None, (1 << 1) | (1 << 9)),
        (2, 5, 7, t0, None),
    ]
    out_ids = torch.empty(16, dtype=torch.int64, device=dev)
    out_sc = torch.empty(16, dtype=torch.float64, device=dev)
    out_cnt = torch.empty(1, dtype=torch.int32, device=dev)
    for qi, user, k, dlo, cmask in calls:
        q = synth.query_vector(spec, qi)
        filt = make_filter(int(hc.offsets[user]), int(hc.offsets[user + 1]), dlo, None, cmask)
        want_ids, want_sc, want_cnt = eng.search_host(q[None, :], [filt], k)
        dq = torch.from_numpy(q).to(dev)
        search_noalloc(0, shard, dq, filt, k, out_ids, out_sc, out_cnt)
        torch.cuda.synchronize()
        n = int(out_cnt.item())
        assert n == int(want_cnt[0]) == k
        assert out_ids[:k].cpu().numpy().view(np.uint64).tolist() == want_ids[0].tolist()
        assert out_sc[:k].cpu().numpy().tolist() == want_sc[0].tolist()
        # ... and what the independent CPU oracle returns for the same call (a parity check in its own right, not only
        # agreement between two routes through the same library)
        oi, os_ = canon.search(hc.vectors, q, k, row_begin=int(hc.offsets[user]), row_end=int(hc.offsets[user + 1]),
                               dates=hc.dates, date_lo=dlo, categories=hc.categories, category_mask=cmask)
        assert out_ids[:k].cpu().numpy().view(np.uint64).tolist() == oi.tolist()
        assert out_sc[:k].cpu().numpy().tolist() == os_.tolist()

    # empty segment: count 0, padded outputs, and the shared scratch is still usable afterwards
    search_noalloc(0, shard, dq, make_filter(0, 0), 5, out_ids, out_sc, out_cnt)
    torch.cuda.synchronize()
    assert int(out_cnt.item()) == 0

    # sizes beyond the stated maximum are refused with a clear message, nothing is launched
    with pytest.raises(_lib.FcbError, match="maximum of 16"):
        search_noalloc(0, shard, dq, filt, 17, out_ids, out_sc, out_cnt)
    big = _lib.FcbShard(vec.data_ptr(), None, None, None, None, None, None, 10, 0, 0, 16384, _lib.FCB_DTYPE_F32)
    with pytest.raises(_lib.FcbError, match="maximum of 8192"):
        search_noalloc(0, big, dq, filt, 5, out_ids, out_sc, out_cnt)
    eng.close()


def test_noalloc_answers_uncertifiable_queries_exactly(built):
    """Many copies of one transaction: the float32 first pass keeps 32 candidates with the same score, so it cannot
    prove its top-k; the engine route then runs a grid-wide float64 kernel with an 8-byte-per-row key buffer. This entry
    point may not allocate one, so its last CTA re-scans the segment itself in float64. Both must return the oracle's
    answer (ties by ascending id), twice in a row."""
    import torch
    from financial_chatbot_llm_b200 import _lib, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter, search_noalloc
    from oracle import canon

    dev = torch.device("cuda", 0)
    spec = synth.SynthSpec(n_rows=5000, dim=256, n_users=2, seed=77, zipf_a=0.0)
    hc = synth.make_host_corpus(spec)
    hc.vectors[100:2400:7] = hc.vectors[50]                      # ~330 exact duplicates inside user 0's segment
    hc.vectors[3000] *= np.float32(1e24)                         # user 1: one row whose norm overflows float32
    vec = torch.from_numpy(hc.vectors).to(dev)
    eng = Engine(device=0, max_batch=1, max_k=16)
    eng.load_shard(vec)
    shard = _lib.FcbShard(vec.data_ptr(), None, None, None, None, None, None, spec.n_rows, 0, 0, spec.dim, _lib.FCB_DTYPE_F32)
    out_ids = torch.empty(16, dtype=torch.int64, device=dev)
    out_sc = torch.empty(16, dtype=torch.float64, device=dev)
    out_cnt = torch.empty(1, dtype=torch.int32, device=dev)
    for qi, user, k in ((0, 0, 10), (1, 0, 16), (2, 1, 5), (3, 1, 12)):
        q = hc.vectors[50].astype(np.float64) + 1e-3 * synth.query_vector(spec, qi) if user == 0 else synth.query_vector(spec, qi)
        filt = make_filter(int(hc.offsets[user]), int(hc.offsets[user + 1]))
        want_ids, want_sc, want_cnt = eng.search_host(q[None, :], [filt], k)
        search_noalloc(0, shard, torch.from_numpy(q).to(dev), filt, k, out_ids, out_sc, out_cnt)
        torch.cuda.synchronize()
        oi, os_ = canon.search(hc.vectors, q, k, row_begin=int(hc.offsets[user]), row_end=int(hc.offsets[user + 1]))
        assert int(out_cnt.item()) == int(want_cnt[0]) == len(oi) == k
        assert out_ids[:k].cpu().numpy().view(np.uint64).tolist() == want_ids[0].tolist() == oi.tolist()
        assert out_sc[:k].cpu().numpy().tolist() == want_sc[0].tolist() == os_.tolist()
    assert eng.exact_reruns == 4, "the engine route must have needed its float64 kernel for all four"
    eng.close()
