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
