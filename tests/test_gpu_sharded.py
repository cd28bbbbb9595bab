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


def test_routed_search_two_ranks_on_one_gpu(built):
    """fcb_search_routed (owner-computes + peer-store exchange) on a 1-GPU box: two engines hold the two point ranges of a
    corpus cut on a user boundary and play the two ranks ONE AFTER THE OTHER on one stream, with the exchange kernel's
    wait half switched off (fcb_debug_exchange_phases(1)): kernels that wait on one another must each have their own
    GPU. The two exchange buffers are ordinary allocations of this process (on a multi-GPU box they are CUDA IPC mappings
    of the peers' memory; the push is the same code). Each rank scans only the queries it owns; after both calls BOTH
    buffers must hold the oracle's answer and the stamp for every query — three times, so both parities are used. The
    wait half runs in the world-1 call at the end (its own stamps) and, between real GPUs, in the CUDA IPC test below."""
    import ctypes as C
    import torch
    from financial_chatbot_llm_b200 import _lib, loader, sharded, synth
    from financial_chatbot_llm_b200.engine import Engine
    from oracle import canon
    lib = _lib.load()
    spec = synth.SynthSpec(n_rows=30000, dim=256, n_users=40, seed=91)
    hc = synth.make_host_corpus(spec, bf16=True)
    world, K, slots = 2, 10, 24
    ranges = sharded.shard_ranges(spec.n_rows, world, hc.offsets)             # the cut lies on a user boundary
    assert ranges[0][1] in set(hc.offsets.tolist())
    cuts = np.array([ranges[0][0], ranges[0][1], ranges[1][1]], dtype=np.int64)
    dev = torch.device("cuda", 0)
    engines = [Engine(device=0, max_batch=slots, max_k=16) for _ in range(world)]
    for e, (lo, hi) in zip(engines, ranges):
        loader.collection_from_host(hc, e, bf16=True, with_ids=False, row_lo=lo, row_hi=hi)
    nbytes = int(lib.fcb_exchange_bytes(slots, K))
    ptrs = []
    for g in range(world):
        p, h = C.c_void_p(), (C.c_ubyte * 64)()
        _lib.check(lib.fcb_peer_alloc(0, nbytes, C.byref(p), h))
        ptrs.append(int(p.value))
    bufs = (C.c_void_p * _lib.FCB_MAX_WORLD)(*([C.c_void_p(p) for p in ptrs] + [None] * (_lib.FCB_MAX_WORLD - world)))
    xs = [_lib.FcbExchange(world, g, bufs, cuts.ctypes.data) for g in range(world)]
    stream = torch.cuda.current_stream(dev).cuda_stream
    status = torch.zeros(world, dtype=torch.int32, device=dev)
    rng = np.random.default_rng(4)
    lib.fcb_debug_exchange_phases(1)                                        # push only: nothing waits for another launch
    for seq in (1, 2, 3):
        users = rng.choice(spec.n_users, size=slots, replace=False)
        users[5] = 0                                                        # an empty filter rides along (count 0)
        farr = sharded.filters_array(hc.offsets[users], hc.offsets[users + 1])
        farr["row_end"][5] = farr["row_begin"][5]
        qs = np.stack([synth.query_vector(spec, 100 * seq + b) for b in range(slots)])
        dq = torch.from_numpy(qs).to(dev)
        torch.cuda.synchronize()
        for g in range(world):
            _lib.check(lib.fcb_search_routed(engines[g]._h, C.byref(xs[g]), dq.data_ptr(), farr.ctypes.data, slots, K, seq,
                                             status[g:].data_ptr(), C.c_void_p(stream)), engines[g]._h)
        torch.cuda.synchronize()
        assert status.cpu().tolist() == [0, 0]
        for g in range(world):
            base = ptrs[g] + (seq & 1) * (nbytes // 2)
            n = slots * K
            ids = torch.as_tensor(sharded._DevicePointer(base, (slots, K), "<i8"), device=dev).cpu().numpy().view(np.uint64)
            sc = torch.as_tensor(sharded._DevicePointer(base + 8 * n, (slots, K), "<f8"), device=dev).cpu().numpy()
            cnt = torch.as_tensor(sharded._DevicePointer(base + 16 * n, (slots,), "<i4"), device=dev).cpu().numpy()
            stamps = torch.as_tensor(sharded._DevicePointer(base + 16 * n + 4 * slots, (slots,), "<i4"), device=dev).cpu().numpy()
            assert stamps.tolist() == [seq] * slots, "every slot of every buffer carries this exchange's stamp"
            for b, u in enumerate(users):
                if b == 5:
                    assert cnt[b] == 0
                    continue
                oi, os_ = canon.search(hc.bf16_bits, qs[b], K, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
                assert cnt[b] == len(oi) and ids[b, :cnt[b]].tolist() == oi.tolist(), (seq, g, b)
                assert sc[b, :cnt[b]].tolist() == os_.tolist(), (seq, g, b)
    # a range that spans both shards is refused (it needs a merge: fcb_search + all-gather + fcb_merge_topk)
    bad = sharded.filters_array(np.full(slots, cuts[1] - 10), np.full(slots, cuts[1] + 10))
    rc = lib.fcb_search_routed(engines[0]._h, C.byref(xs[0]), dq.data_ptr(), bad.ctypes.data, slots, K, 9, status.data_ptr(),
                               C.c_void_p(stream))
    assert rc == _lib.FCB_ERR_UNSUPPORTED
    lib.fcb_debug_exchange_phases(3)
    for e in engines:
        e.close()
    for p in ptrs:
        lib.fcb_peer_close(0, C.c_void_p(p), 1)
    # world of one rank, push AND wait (the production kernel): the rank owns everything and waits for its own stamps
    eng = Engine(device=0, max_batch=slots, max_k=16)
    col = loader.collection_from_host(hc, eng, bf16=True, with_ids=False)
    xch = sharded.PeerExchange(0, 1, 0, slots, K, [0, spec.n_rows])
    routed = sharded.RoutedSearcher(col, xch)
    for it in range(2):
        users = rng.choice(spec.n_users, size=slots, replace=False)
        qs = np.stack([synth.query_vector(spec, 900 + 10 * it + b) for b in range(slots)])
        ids, sc, cnt = routed.search(qs, sharded.filters_array(hc.offsets[users], hc.offsets[users + 1]), K)
        for b, u in enumerate(users):
            oi, os_ = canon.search(hc.bf16_bits, qs[b], K, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
            assert ids[b, :cnt[b]].tolist() == oi.tolist() and sc[b, :cnt[b]].tolist() == os_.tolist()
    xch.close()
    eng.close()


def routed_worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from financial_chatbot_llm_b200 import loader, sharded, synth
        from financial_chatbot_llm_b200.engine import Engine
        from oracle import canon
        spec = synth.SynthSpec(n_rows=40000, dim=256, n_users=60, seed=93)
        hc = synth.make_host_corpus(spec)
        ranges = sharded.shard_ranges(spec.n_rows, world, hc.offsets)
        lo, hi = ranges[rank]
        slots, K = 16 * world, 10
        eng = Engine(device=rank, max_batch=slots, max_k=16)
        col = loader.collection_from_host(hc, eng, with_ids=False, row_lo=lo, row_hi=hi)
        xch = sharded.PeerExchange(rank, world, rank, slots, K, [r[0] for r in ranges] + [ranges[-1][1]])
        routed = sharded.RoutedSearcher(col, xch)
        rng = np.random.default_rng(5)
        ok = True
        for it in range(3):
            users = rng.choice(spec.n_users, size=slots, replace=False)
            qs = np.stack([synth.query_vector(spec, 10 * it + b) for b in range(slots)])
            ids, sc, cnt = routed.search(qs, sharded.filters_array(hc.offsets[users], hc.offsets[users + 1]), K)
            for b, u in enumerate(users):
                oi, os_ = canon.search(hc.vectors, qs[b], K, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
                ok &= ids[b, :cnt[b]].tolist() == oi.tolist() and sc[b, :cnt[b]].tolist() == os_.tolist()
        ret[rank] = bool(ok)
        dist.barrier()
        xch.close()
        eng.close()
    finally:
        dist.destroy_process_group()


def test_routed_search_over_cuda_ipc_matches_oracle(built):
    """The same over real peer mappings: one process per GPU, exchange buffers opened through CUDA IPC (>= 2 GPUs)."""
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    port = 29700 + (os.getpid() % 2000)
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(routed_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {r: True for r in range(world)}
