"""Part of bench.py (bench infrastructure, not product code): BASELINE.json configs[3] for bench.py --gpus N (N >= 2): 100M x 1024-d bf16 sharded by point range over the N GPUs of
one box, 100 000 tenants (user_id filter), batches of 64 distinct tenants, top-10. Three measurements, each timed with CUDA
events (max over ranks) after warm-up, inputs resident in HBM:

  tenants_closed_loop  ONE batch of 64 in flight: every rank scans the batch against its shard, one NCCL all-gather of the
                       per-shard candidates, one k-way merge launch. Three serialized latency-bound steps: the time does not
                       depend on N (this is what round 1 measured).
  tenants_routed       owner-computes: shard cuts lie on user boundaries, so a tenant lives on ONE GPU. A step carries N batches
                       of 64 (one per front-end rank, N x 64 queries); every rank scans only the queries it owns (~64) and
                       writes their final top-10 straight into every rank's exchange buffer over NVLink peer mappings
                       (fcb_search_routed: plain stores + a stamp, then a bounded wait for everybody else's). No NCCL call,
                       no merge launch. Per-rank work per step is constant, so queries/s grows with N. Measured with one
                       engine per GPU (N batches in flight) and with two engines per GPU alternating steps on two streams
                       (2N in flight: the exchange wait of one step overlaps the scan of the next).
  whale_striped        one query whose segment is the whole corpus: N/G rows per GPU, all-gather + merge: bandwidth scaling.
"""
from __future__ import annotations

import os
import time

import numpy as np

ROWS, DIM, TENANTS, BATCH, K = 100_000_000, 1024, 100_000, 64, 10


def run(dev, world: int, rank: int, steps: int, timed_steps, rows: int = ROWS, tenants: int = TENANTS):
    import torch
    import torch.distributed as dist
    from financial_chatbot_llm_b200 import loader, sharded, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon                                          # checker only (parity flags below)

    spec = synth.SynthSpec(n_rows=rows, dim=DIM, seed=44, n_users=tenants, zipf_a=0.0)   # ~equal tenants (~1000 rows)
    offsets = synth.segment_offsets(spec)
    ranges = sharded.shard_ranges(rows, world, offsets)               # cuts on user boundaries: a tenant never straddles
    lo, hi = ranges[rank]
    slots = BATCH * world
    eng = Engine(device=dev.index, max_batch=slots, max_k=16)
    t0 = time.perf_counter()
    col = loader.collection_synthetic(spec, eng, bf16=True, row_lo=lo, row_hi=hi, with_cents=False)
    torch.cuda.synchronize()
    gen_s = time.perf_counter() - t0
    rng = np.random.default_rng(7)
    n_sets = 32

    def check(ids, sc, cnt, tenant_list, qbase, which):
        ok = True
        for b in which:
            t = int(tenant_list[b]); a, e_ = int(offsets[t]), int(offsets[t + 1])
            bits = synth.vectors_bf16_bits(spec, a, e_).reshape(e_ - a, DIM)
            oi_, os_ = canon.search(bits, synth.query_vector(spec, qbase + b), K, id_base=a)
            ok = ok and ids[b, :cnt[b]].tolist() == oi_.tolist() and sc[b, :cnt[b]].tolist() == os_.tolist()
        return ok

    def all_ok(flag: bool) -> bool:
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    # ---- closed loop: one batch of 64, all-gather + merge
    searcher = sharded.ShardedSearcher(col, world, dev)
    tsets = [rng.choice(tenants, size=BATCH, replace=False) for _ in range(n_sets)]
    fsets = [[make_filter(int(offsets[t]), int(offsets[t + 1])) for t in ts] for ts in tsets]
    qsets = [torch.from_numpy(np.stack([synth.query_vector(spec, s * 1000 + b) for b in range(BATCH)])).to(dev) for s in range(4)]
    m = searcher.merger(BATCH, K)
    oi, osc, oc = m.local_views

    def step_closed(i):
        eng.search_device(qsets[i % 4], fsets[i % n_sets], K, oi, osc, oc)
        m.merge()
    ms_c, per_c, _ = timed_steps(step_closed, steps, 5, dev, world)
    step_closed(0)
    torch.cuda.synchronize()
    ok_closed = all_ok(check(m.ids.cpu().numpy().view(np.uint64), m.scores.cpu().numpy(), m.counts.cpu().numpy(), tsets[0], 0, range(3)))

    # ---- routed: N batches of 64 in flight per engine, owner-computes + peer-store exchange. TWO engines per GPU share the
    #      shard's columns (each has its own scratch, exchange buffers and stream) and alternate steps, so the exchange wait of
    #      one step overlaps the scan of the next: 2N batches in flight.
    cuts = [r[0] for r in ranges] + [ranges[-1][1]]
    routed_res = None
    try:
        engines, exchanges, searchers, streams = [eng], [], [], []
        n_engines = max(2, int(os.environ.get("FCB_CFG4_ENGINES", "2")))
        vec, _ids, _rank, dts, cats, _cents, offs_d = eng._keep
        for _ in range(n_engines - 1):
            e_more = Engine(device=dev.index, max_batch=slots, max_k=16)
            e_more.load_shard(vec, dates=dts, categories=cats, seg_offsets=offs_d, row_base=lo)
            engines.append(e_more)
        from financial_chatbot_llm_b200.client import Collection
        for e in engines:
            xch = sharded.PeerExchange(dev.index, world, rank, slots, K, cuts)
            exchanges.append(xch)
            searchers.append(sharded.RoutedSearcher(Collection("transactions", e, col.users, col.offsets, col.payload_fn,
                                                               n_total_rows=rows), xch))
            streams.append(torch.cuda.Stream(device=dev))
        rsets = [rng.choice(tenants, size=slots, replace=False) for _ in range(n_sets)]
        farr = [sharded.filters_array(offsets[ts], offsets[ts + 1]) for ts in rsets]
        rq = [torch.from_numpy(np.stack([synth.query_vector(spec, 50_000 + s * 1000 + b) for b in range(slots)])).to(dev) for s in range(2)]
        torch.cuda.synchronize()

        def run_routed(n_eng, n_steps, warm):
            """n_steps steps alternating over n_eng engines / streams; CUDA events on each stream; returns ms per step (max over
            ranks), this rank's sorted per-step spans, clocks."""
            for i in range(warm):
                searchers[i % n_eng].search_device(rq[i % 2], farr[i % n_sets], K, stream=streams[i % n_eng].cuda_stream)
            torch.cuda.synchronize()
            dist.barrier()
            from bench import ClockSampler
            sampler = ClockSampler(dev.index).start() if rank == 0 else None
            time.sleep(0.3 if rank == 0 else 0.0)
            dist.barrier()
            torch.cuda.synchronize()
            t_a = time.perf_counter()
            first = [torch.cuda.Event(enable_timing=True) for _ in range(n_eng)]
            last = [torch.cuda.Event(enable_timing=True) for _ in range(n_eng)]
            marks = []
            for j in range(n_eng):
                first[j].record(streams[j])
            for i in range(n_steps):
                j = i % n_eng
                searchers[j].search_device(rq[i % 2], farr[i % n_sets], K, stream=streams[j].cuda_stream)
                ev = torch.cuda.Event(enable_timing=True)
                ev.record(streams[j])
                marks.append(ev)
            for j in range(n_eng):
                last[j].record(streams[j])
            torch.cuda.synchronize()
            dist.barrier()
            t_b = time.perf_counter()
            span = max(first[0].elapsed_time(last[j]) for j in range(n_eng))          # streams start together (synchronised above)
            tm = torch.tensor([span], dtype=torch.float64, device=dev)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            per = sorted(marks[i].elapsed_time(marks[i + n_eng]) / n_eng for i in range(len(marks) - n_eng))
            return float(tm.item()) / n_steps, per, (sampler.stop(t_a, t_b) if sampler else None)

        def run_routed_threads(n_eng, n_steps, warm):
            """The same steps issued by one host thread PER ENGINE (ctypes releases the GIL inside the library call), so the
            ~50 us of host work per call overlap; ms per step = the longest stream's span / n_steps, max over ranks."""
            import threading
            per_eng = n_steps // n_eng
            for i in range(warm):
                searchers[i % n_eng].search_device(rq[i % 2], farr[i % n_sets], K, stream=streams[i % n_eng].cuda_stream)
            torch.cuda.synchronize()
            dist.barrier()
            first = [torch.cuda.Event(enable_timing=True) for _ in range(n_eng)]
            last = [torch.cuda.Event(enable_timing=True) for _ in range(n_eng)]
            errors = []
            go = threading.Barrier(n_eng)

            def worker(j):
                try:
                    torch.cuda.set_device(dev)
                    go.wait()
                    first[j].record(streams[j])
                    for i in range(per_eng):
                        searchers[j].search_device(rq[i % 2], farr[(i * n_eng + j) % n_sets], K, stream=streams[j].cuda_stream)
                    last[j].record(streams[j])
                except Exception as exc:                                  # surfaced below: a failed thread must not hang the ranks
                    errors.append(repr(exc))
            threads = [threading.Thread(target=worker, args=(j,)) for j in range(n_eng)]
            for t in threads:
                t.start()
            for t in threads:
                t.join()
            torch.cuda.synchronize()
            dist.barrier()
            if errors:
                raise RuntimeError("; ".join(errors))
            span = max(first[i].elapsed_time(last[j]) for i in range(n_eng) for j in range(n_eng))
            tm = torch.tensor([span], dtype=torch.float64, device=dev)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            return float(tm.item()) / (per_eng * n_eng)

        def check_routed(j):
            seq = searchers[j].search_device(rq[0], farr[0], K, stream=streams[j].cuda_stream)
            torch.cuda.synchronize()
            ids, sc, cnt = exchanges[j].results(seq)
            ids, sc, cnt = ids.cpu().numpy().view(np.uint64), sc.cpu().numpy(), cnt.cpu().numpy()
            # every rank checks slots owned by DIFFERENT ranks in its own buffer (so the delivery itself is checked)
            which = [rank, (rank * BATCH + 7) % slots, slots - 1 - rank, slots // 2]
            return int(exchanges[j].status.item()) == 0 and bool((cnt == K).all()) and check(ids, sc, cnt, rsets[0], 50_000, which)

        n_r = max(steps, 40)
        ms_1, per_1, _ = run_routed(1, n_r, 6)
        ms_2, per_2, clocks = run_routed(2, n_r, 6)
        more = {}
        for ne in range(3, n_engines + 1):                              # experiments: FCB_CFG4_ENGINES > 2
            ms_n, per_n, _ = run_routed(ne, n_r, 6)
            more[f"{ne}_engines_per_gpu"] = {"ms_per_step": ms_n, "p50_ms": per_n[len(per_n) // 2], "queries_per_s": slots * 1e3 / ms_n}
        if os.environ.get("FCB_CFG4_THREADS"):                           # experiment: one issuing thread per engine
            for ne in range(2, n_engines + 1):
                ms_t = run_routed_threads(ne, n_r * 2, 6)
                more[f"{ne}_engines_per_gpu_one_thread_each"] = {"ms_per_step": ms_t, "queries_per_s": slots * 1e3 / ms_t}
        ok_routed = all_ok(all(check_routed(j) for j in range(n_engines)))
        owned = int(np.sum((offsets[rsets[0]] >= lo) & (offsets[rsets[0]] < hi)))
        routed_res = {"ms_per_step": ms_2, "p50_ms": per_2[len(per_2) // 2], "queries_per_step": slots, "engines_per_gpu": 2,
                      "batches_in_flight": 2 * world, "queries_per_s": slots * 1e3 / ms_2,
                      "one_engine_per_gpu": {"ms_per_step": ms_1, "p50_ms": per_1[len(per_1) // 2], "batches_in_flight": world,
                                             "queries_per_s": slots * 1e3 / ms_1},
                      "queries_owned_by_rank0_in_a_step": owned, "launches_per_step_per_rank": eng.last_search_launches,
                      "nccl_calls_per_step": 0, "parity": ok_routed, "clocks": clocks,
                      "exact_reruns": sum(e.exact_reruns for e in engines), **more}
        dist.barrier()
        for xch in exchanges:
            xch.close()
        for e_more in engines[1:]:
            e_more.close()
    except Exception as e:                                            # e.g. CUDA IPC unavailable on this box: the closed loop still stands
        import traceback
        routed_res = {"error": f"{type(e).__name__}: {e}", "trace": traceback.format_exc()[-600:]}

    # ---- whale: one query over the whole corpus, striped over all shards
    qw = torch.from_numpy(synth.query_vector(spec, 999)[None, :]).to(dev)
    m1 = searcher.merger(1, K)
    wi, ws, wc = m1.local_views
    fw = [make_filter(0, rows)]

    def step_whale(i):
        eng.search_device(qw, fw, K, wi, ws, wc)
        m1.merge()
    ms_w, per_w, _ = timed_steps(step_whale, max(20, steps // 4), 3, dev, world)
    wid, wsc = m1.ids.cpu().numpy().view(np.uint64)[0], m1.scores.cpu().numpy()[0]
    ok_whale = bool(np.all(np.diff(wsc) <= 0))
    for j in (0, K - 1):                                              # returned rows re-scored from host-regenerated rows
        r = int(wid[j])
        ok_whale = ok_whale and canon.scores(synth.vectors_bf16_bits(spec, r, r + 1).reshape(1, DIM), synth.query_vector(spec, 999))[0] == wsc[j]
    ok_whale = all_ok(ok_whale)
    gbs = rows * DIM * 2 / (ms_w / 1e3) / 1e9
    seg_rows = float(np.mean(np.diff(offsets)))
    eng.close()
    del eng, col
    torch.cuda.empty_cache()
    return {
        "workload": f"configs[3]: {rows} x {DIM} bf16 sharded by point range over {world} GPUs (cuts on user boundaries), {tenants} tenants "
                    f"(~{seg_rows:.0f} rows each), user_id filter, top-{K}, batches of {BATCH} distinct tenants",
        "rows_per_gpu": hi - lo, "corpus_gen_s": gen_s,
        "tenants_closed_loop": {"ms_per_batch": ms_c, "p50_ms": per_c[len(per_c) // 2], "queries_per_s": BATCH * 1e3 / ms_c,
                                "batches_in_flight": 1, "nccl_calls_per_step": 1, "parity": ok_closed,
                                "how": "fcb_search over the whole batch on every rank + all_gather_into_tensor + fcb_merge_topk"},
        "tenants_routed": routed_res,
        "whale_striped": {"ms_per_query": ms_w, "p50_ms": per_w[len(per_w) // 2], "aggregate_GBs": gbs, "per_gpu_GBs": gbs / world, "parity": ok_whale},
        "algorithmic_bytes_per_query": seg_rows * DIM * 2,
    }
