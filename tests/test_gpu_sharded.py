"""Sharded search over NCCL on >= 2 GPUs (skipped on a 1-GPU box; the host logic is covered on CPU by
tests/test_sharded_host.py). Tenants straddle the shard cut; every rank must end with the single-shard
oracle answer, bit-exact."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from financial_chatbot_llm_b200 import loader, sharded, synth
        from financial_chatbot_llm_b200.engine import Engine, make_filter
        from oracle import canon
        spec = synth.SynthSpec(n_rows=24000, dim=256, n_users=9, seed=77)
        hc = synth.make_host_corpus(spec)
        lo, hi = sharded.shard_ranges(spec.n_rows, world)[rank]
        eng = Engine(device=rank, max_batch=8, max_k=64)
        col = loader.collection_from_host(hc, eng, row_lo=lo, row_hi=hi)
        s = sharded.ShardedSearcher(col, world, dev)
        t0 = int(np.median(hc.dates))
        ok = True
        for k in (10, 40):
            qs = np.stack([synth.query_vector(spec, 10 * k + i) for i in range(6)])
            users = [0, 1, 2, 4, 8, 3]
            fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]), t0 if i % 2 else None) for i, u in enumerate(users)]
            ids, sc, cnt = s.search(qs, fl, k)
            for i, u in enumerate(users):
                oi, os_ = canon.search(hc.vectors, qs[i], k, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]),
                                       ids=hc.ids, dates=hc.dates, date_lo=t0 if i % 2 else None)
                ok &= ids[i, :cnt[i]].tolist() == oi.tolist() and sc[i, :cnt[i]].tolist() == os_.tolist()
        ret[rank] = bool(ok)
        eng.close()
    finally:
        dist.destroy_process_group()


def test_nccl_sharded_search_matches_oracle(built):
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    port = 29600 + (os.getpid() % 2000)
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {r: True for r in range(world)}


def test_two_shards_on_one_gpu_client_payloads(built):
    """The sharded client on a 1-GPU box: two engines hold the two point ranges, their packed results are laid out
    as the all-gather would and merged by fcb_merge_topk; query_points must return the oracle's points WITH the
    payload of each point's own row — including points that live on the other shard (permuted ids, so id != row)."""
    import torch
    from financial_chatbot_llm_b200 import loader, sharded, synth
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200.engine import Engine
    from oracle import canon
    spec = synth.SynthSpec(n_rows=20000, dim=128, n_users=6, seed=5)
    hc = synth.make_host_corpus(spec)
    rng = np.random.default_rng(3)
    hc.ids = rng.permutation(spec.n_rows).astype(np.uint64) * 3 + 11          # unique, not ascending, id != row
    dev = torch.device("cuda", 0)
    world = 2
    ranges = sharded.shard_ranges(spec.n_rows, world)
    engines = [Engine(device=0, max_batch=4, max_k=32) for _ in range(world)]
    cols = [loader.collection_from_host(hc, e, row_lo=lo, row_hi=hi) for e, (lo, hi) in zip(engines, ranges)]

    class TwoShardSearch:
        def search(self, queries, filters, k):
            B = queries.shape[0]
            mg = sharded.ShardMerger(dev, world, B, k)
            dq = torch.from_numpy(np.ascontiguousarray(queries)).to(dev)
            for g, e in enumerate(engines):
                ids, sc, cnt = sharded.ShardMerger._views(mg.gathered[g], B, k)
                e.search_device(dq, list(filters), k, ids, sc, cnt)
            sharded.cuda_merge(0, mg.gathered, world, B, k, mg.merged)
            return mg.ids.cpu().numpy().view(np.uint64), mg.scores.cpu().numpy(), mg.counts.cpu().numpy()

    su = int(np.searchsorted(hc.offsets, ranges[0][1], side="right") - 1)     # the user straddling the cut
    assert hc.offsets[su] < ranges[0][1] < hc.offsets[su + 1]
    for rank in range(world):                                                 # "every rank makes the same call"
        cl = sharded.ShardedB200Client(world, dev)
        cl.attach_collection(cols[rank])
        cl._searchers["transactions"] = TwoShardSearch()
        for u, k in ((su, 12), (0, 5), (5, 20)):
            q = synth.query_vector(spec, 100 + u)
            pts = cl.query_points("transactions", query=q.tolist(), limit=k, query_filter=m.Filter(must=[
                m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[u]))])).points
            oi, os_ = canon.search(hc.vectors, q, k, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]), ids=hc.ids)
            assert [p.id for p in pts] == oi.tolist() and [p.score for p in pts] == os_.tolist()
            row_of = {int(i): r for r, i in enumerate(hc.ids)}
            for p in pts:
                assert p.payload == hc.payload(row_of[p.id]), "payload of the point's own row, whichever shard holds it"
                assert p.payload["metadata"]["user_id"] == hc.users[u]
    for e in engines:
        e.close()
