"""Host-side logic of the sharded path on CPU: world_size-2 gloo process group, fake per-shard
engines (numpy scoring through the ORACLE, test only) and a numpy merge injected in place of the
CUDA merge kernel. Checks the packed buffer layout, the all-gather, shard clipping and that every
rank ends with the single-shard answer."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def numpy_merge(gathered, world, batch, k, merged):
    from financial_chatbot_llm_b200.sharded import ShardMerger
    outs = ShardMerger._views(merged, batch, k)
    for b in range(batch):
        cand = []
        for g in range(world):
            ids, sc, cnt = ShardMerger._views(gathered[g], batch, k)
            for j in range(int(cnt[b])):
                cand.append((-float(sc[b, j]), int(ids[b, j]) & (2 ** 64 - 1)))
        cand.sort()
        n = min(k, len(cand))
        outs[2][b] = n
        for j in range(k):
            outs[0][b, j] = int(np.array(cand[j][1] if j < n else 2 ** 64 - 1, dtype=np.uint64).view(np.int64))
            outs[1][b, j] = -cand[j][0] if j < n else -np.inf


class FakeEngine:
    """Per-shard exact search through the C oracle (rows [row_lo,row_hi) of the host corpus)."""
    def __init__(self, hc, row_lo, row_hi):
        self.hc, self.row_base, self.n = hc, row_lo, row_hi - row_lo
        self.dim = hc.spec.dim

    def search_device(self, dq, filters, k, ids, scores, counts, stream=None):
        from oracle import canon
        lo, hi = self.row_base, self.row_base + self.n
        for b, f in enumerate(filters):
            rb, re_ = max(f.row_begin, lo), min(f.row_end, hi)
            oi, os_ = (np.zeros(0, np.uint64), np.zeros(0)) if re_ <= rb else canon.search(
                self.hc.vectors, dq[b].numpy(), k, row_begin=rb, row_end=re_, ids=self.hc.ids,
                dates=self.hc.dates, date_lo=None if f.date_lo < -2 ** 62 else f.date_lo,
                date_hi=None if f.date_hi > 2 ** 62 else f.date_hi)
            counts[b] = len(oi)
            ids[b, :len(oi)] = torch.from_numpy(oi.view(np.int64))
            scores[b, :len(oi)] = torch.from_numpy(os_)


def worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from financial_chatbot_llm_b200 import qdrant_models as m
        from financial_chatbot_llm_b200 import sharded, synth
        from financial_chatbot_llm_b200.client import Collection
        from financial_chatbot_llm_b200.engine import make_filter
        from oracle import canon
        spec = synth.SynthSpec(n_rows=3000, dim=64, n_users=7, seed=21)
        hc = synth.make_host_corpus(spec)
        ranges = sharded.shard_ranges(spec.n_rows, world)
        lo, hi = ranges[rank]
        eng = FakeEngine(hc, lo, hi)
        col = Collection("transactions", eng, hc.users, hc.offsets, hc.payload, n_total_rows=spec.n_rows)
        dev = torch.device("cpu")
        s = sharded.ShardedSearcher(col, world, dev, merge_impl=numpy_merge)
        t0 = int(np.median(hc.dates))
        qs = np.stack([synth.query_vector(spec, i) for i in range(5)])
        su = int(np.searchsorted(hc.offsets, ranges[0][1], side="right") - 1)   # the user straddling the cut
        users = [0, su, 6, 2, 5]
        fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]), t0 if i % 2 else None) for i, u in enumerate(users)]
        ids, sc, cnt = s.search(qs, fl, 12)
        ok = True
        for i, u in enumerate(users):
            oi, os_ = canon.search(hc.vectors, qs[i], 12, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]),
                                   ids=hc.ids, dates=hc.dates, date_lo=t0 if i % 2 else None)
            ok &= ids[i, :cnt[i]].tolist() == oi.tolist() and sc[i, :cnt[i]].tolist() == os_.tolist()
        straddles = any(hc.offsets[u] < ranges[0][1] < hc.offsets[u + 1] for u in users)
        # the qdrant-shaped collective call
        cl = sharded.ShardedB200Client(world, dev, merge_impl=numpy_merge)
        cl.attach_collection(col)
        pts = cl.query_points("transactions", query=qs[1].tolist(), limit=5, query_filter=m.Filter(must=[
            m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[3]))])).points
        oi, _ = canon.search(hc.vectors, qs[1], 5, row_begin=int(hc.offsets[3]), row_end=int(hc.offsets[4]))
        ok &= [p.id for p in pts] == oi.tolist() and all(p.payload["metadata"]["user_id"] == hc.users[3] for p in pts)
        # OR / NOT filters: several conjuncts ride in one collective search and are merged on the host (client.py)
        from oracle import local_mode
        flt = m.Filter(should=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[u])) for u in (1, su, 5)],
                       must_not=[m.FieldCondition(key="metadata.date", range=m.Range(gte=t0 - 40 * 86400, lte=t0 + 40 * 86400))])
        pts = cl.query_points("transactions", query=qs[2].tolist(), limit=9, query_filter=flt).points
        mask = local_mode.payload_mask([hc.payload(r) for r in range(spec.n_rows)], flt)
        rows = np.flatnonzero(mask)
        want = sorted(((-float(canon.scores(hc.vectors[r:r + 1], qs[2])[0]), int(hc.ids[r])) for r in rows))[:9]
        ok &= [p.id for p in pts] == [w[1] for w in want] and len(pts) == 9
        ret[rank] = (bool(ok), bool(straddles))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_sharded_search(built):
    world = 2
    port = 29500 + (os.getpid() % 2000)
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {0: (True, True), 1: (True, True)}


def test_shard_ranges_cover_and_align():
    from financial_chatbot_llm_b200 import sharded, synth
    assert sharded.shard_ranges(10, 4) == [(0, 2), (2, 5), (5, 8), (8, 10)]
    off = synth.segment_offsets(synth.SynthSpec(n_rows=100000, n_users=1000))
    r = sharded.shard_ranges(100000, 8, off)
    assert r[0][0] == 0 and r[-1][1] == 100000 and all(a[1] == b[0] for a, b in zip(r, r[1:]))
    assert all(c in set(off.tolist()) for _, c in r)
    assert sharded.packed_len(64, 10) == 1312


def test_row_of_id_global_map_and_miss():
    """A sharded answer names points of every shard: the id -> row map must be global, and an id it does not hold
    raises (round-1 fault: searchsorted's neighbour was returned, i.e. another transaction's payload)."""
    from financial_chatbot_llm_b200.client import Collection

    class Eng:
        row_base, dim = 1000, 8
    ids_all = (np.arange(3000, dtype=np.uint64) * 7 + 5)[::-1].copy()          # descending: row != id, rank != row
    users, offsets = ["u0", "u1"], np.array([0, 1500, 3000])
    col = Collection("t", Eng(), users, offsets, lambda row: {"row": row}, n_total_rows=3000,
                     ids_host=ids_all, ids_row_base=0)
    for row in (0, 999, 1000, 1999, 2000, 2999):
        assert col.row_of_id(int(ids_all[row])) == row
    with pytest.raises(KeyError):
        col.row_of_id(6)                                                       # between two ids
    with pytest.raises(KeyError):
        col.row_of_id(10 ** 9)                                                 # beyond the last id
    # a map over the local slice only: local ids resolve, ids of other shards raise instead of aliasing
    loc = Collection("t", Eng(), users, offsets, lambda row: {"row": row}, n_total_rows=3000, ids_host=ids_all[1000:2000])
    assert loc.row_of_id(int(ids_all[1500])) == 1500
    with pytest.raises(KeyError):
        loc.row_of_id(int(ids_all[5]))
    # ids == row numbers (no id column): range-checked
    plain = Collection("t", Eng(), users, offsets, lambda row: {"row": row}, n_total_rows=3000)
    assert plain.row_of_id(2999) == 2999
    with pytest.raises(KeyError):
        plain.row_of_id(3000)


def test_filters_array_has_the_layout_of_fcb_filter():
    """Batches of filters travel to fcb_search_routed as ONE numpy structured array: same bytes as an FcbFilter array."""
    import ctypes as C
    from financial_chatbot_llm_b200 import _lib, sharded
    from financial_chatbot_llm_b200.engine import make_filter
    rb, re_ = np.array([0, 10, 500]), np.array([10, 25, 500])
    arr = sharded.filters_array(rb, re_)
    assert arr.dtype.itemsize == C.sizeof(_lib.FcbFilter) == 40
    want = (_lib.FcbFilter * 3)(*[make_filter(int(a), int(b)) for a, b in zip(rb, re_)])
    assert arr.tobytes() == bytes(want)
    arr = sharded.filters_array(rb, re_, date_lo=5, date_hi=9, category_mask=3)
    want = (_lib.FcbFilter * 3)(*[make_filter(int(a), int(b), 5, 9, 3) for a, b in zip(rb, re_)])
    assert arr.tobytes() == bytes(want)
