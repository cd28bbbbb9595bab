#!/usr/bin/env python
"""Secondary measurements (one JSON line each) beside bench.py's headline:
   cfg2a-50   10M x 1536 fp32 whale, date filter passing ~50 %
   cfg2b      10M x 1536 fp32, 10 000 Zipf tenants, batch 1 random tenants (latency-bound)
   cfg2b-B64  same corpus, 64 random tenants per call (grouped launch)
   k10000     whale, k = 10000 (general-k path)
   bf16-b1    10M x 1536 bf16 whale, batch 1
   cfg5       1B-row aggregation (user x category x month), 100k users
   fused      whale, k = 10000 retrieval + (category, month) group-by of the returned points, both on the device
Timing: CUDA events on the launching stream, >= 3 warm-ups, inputs >> L2 (or stated)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import __graft_entry__ as graft
from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter, synth_columns

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
DEV = torch.device("cuda", 0)


def timed(fn, steps, warmup=3):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        fn(i)
        evs[i + 1].record()
    torch.cuda.synchronize()
    per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
    return evs[0].elapsed_time(evs[-1]) / steps, per[len(per) // 2]


def emit(name, ms, p50, alg_bytes, **extra):
    gbs = alg_bytes / (ms / 1e3) / 1e9
    print(json.dumps({"workload": name, "ms": ms, "p50_ms": p50, "algorithmic_bytes": alg_bytes, "achieved_GBs": gbs,
                      "frac_measured_peak": gbs / PEAK, "frac_8TBs": gbs / 8000.0, **extra}), flush=True)


def search_bench(args):
    rows, dim, K = args.rows, 1536, 10
    # ---- whale corpus fp32
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=1, n_users=1)
    eng = Engine(device=0, max_batch=64, max_k=10000)
    loader.collection_synthetic(spec, eng, with_cents=False)
    qs = [torch.from_numpy(synth.query_vector(spec, i)[None, :]).to(DEV) for i in range(8)]
    oi = torch.empty((64, 10000), dtype=torch.int64, device=DEV)
    os_ = torch.empty((64, 10000), dtype=torch.float64, device=DEV)
    oc = torch.empty(64, dtype=torch.int32, device=DEV)
    t_mid = synth.DATE_BASE + spec.date_span_seconds // 2
    f50 = [make_filter(0, rows, t_mid, None)]
    ms, p50 = timed(lambda i: eng.search_device(qs[i % 8], f50, K, oi, os_, oc), args.steps)
    n_pass = int((eng._keep[3] >= t_mid).sum().item())
    emit("cfg2a-50: 10M x 1536 fp32 whale, date>=t_mid", ms, p50, n_pass * dim * 4 + rows * 4, rows_pass=n_pass, qps=1e3 / ms)
    fall = [make_filter(0, rows, synth.DATE_BASE, None)]
    ms, p50 = timed(lambda i: eng.search_device(qs[i % 8], fall, 10000, oi, os_, oc), args.steps)
    emit("k10000: whale, k=10000 general path", ms, p50, rows * dim * 4 + rows * 4 + rows * 8, qps=1e3 / ms,
         launches=eng.last_search_launches)
    ms, p50 = timed(lambda i: eng.search_device(qs[i % 8], fall, 100, oi, os_, oc), args.steps)
    emit("k100: whale, k=100 general path", ms, p50, rows * dim * 4 + rows * 4 + rows * 8, qps=1e3 / ms)
    eng.close(); del eng
    torch.cuda.empty_cache()

    # ---- multi-tenant fp32
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=2, n_users=10000, zipf_a=1.1)
    eng = Engine(device=0, max_batch=64, max_k=16)
    col = loader.collection_synthetic(spec, eng, with_cents=False)
    off = col.offsets
    rng = np.random.default_rng(0)
    tenants = rng.integers(0, spec.n_users, size=4096)
    fl = [[make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)] for t in tenants]
    segb = np.array([int(off[t + 1] - off[t]) for t in tenants], dtype=np.int64)
    ms, p50 = timed(lambda i: eng.search_device(qs[i % 8], fl[i % 4096], K, oi, os_, oc), 2000, warmup=20)
    emit("cfg2b: 10k Zipf tenants, batch 1, uniform random tenant", ms, p50, float(segb[:2000].mean()) * (dim * 4 + 4),
         qps=1e3 / ms, mean_segment_rows=float(segb[:2000].mean()), median_segment_rows=float(np.median(segb)))
    q64 = torch.from_numpy(np.stack([synth.query_vector(spec, 100 + i) for i in range(64)])).to(DEV)
    fl64 = [[fl[(s * 64 + j) % 4096][0] for j in range(64)] for s in range(64)]
    ms, p50 = timed(lambda i: eng.search_device(q64, fl64[i % 64], K, oi, os_, oc), 300, warmup=10)
    emit("cfg2b-B64: 64 random tenants per call", ms, p50, float(segb.mean()) * 64 * (dim * 4 + 4), qps=64e3 / ms)
    # host-API latency (wall clock, includes H2D/D2H + sync)
    qh = synth.query_vector(spec, 5)[None, :]
    for i in range(20):
        eng.search_host(qh, fl[i], K)
    t = time.perf_counter()
    lat = []
    for i in range(1000):
        t1 = time.perf_counter(); eng.search_host(qh, fl[i], K); lat.append(time.perf_counter() - t1)
    print(json.dumps({"workload": "cfg2b host API (fcb_search_host) wall clock", "qps": 1000 / (time.perf_counter() - t),
                      "p50_us": 1e6 * float(np.median(lat)), "p99_us": 1e6 * float(np.percentile(lat, 99))}), flush=True)
    eng.close(); del eng, col
    torch.cuda.empty_cache()

    # ---- bf16 whale batch 1
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=3, n_users=1)
    eng = Engine(device=0, max_batch=4, max_k=16)
    loader.collection_synthetic(spec, eng, bf16=True, with_cents=False)
    ms, p50 = timed(lambda i: eng.search_device(qs[i % 8], fall, K, oi, os_, oc), args.steps)
    emit("bf16-b1: 10M x 1536 bf16 whale, batch 1", ms, p50, rows * dim * 2 + rows * 4, qps=1e3 / ms)
    eng.close(); del eng
    torch.cuda.empty_cache()

    # ---- cfg3: bf16, 256 concurrent queries on one shared segment (tcgen05 path)
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=3, n_users=1)
    eng = Engine(device=0, max_batch=256, max_k=16)
    loader.collection_synthetic(spec, eng, bf16=True, with_cents=False)
    from bench import ClockSampler
    for B in (256, 128, 64):
        qB = torch.from_numpy(np.stack([synth.query_vector(spec, 500 + i) for i in range(B)])).to(DEV)
        fB = [make_filter(0, rows)] * B
        sampler = ClockSampler(0).start()
        time.sleep(0.2)
        t_a = time.perf_counter()
        ms, p50 = timed(lambda i: eng.search_device(qB, fB, K, oi, os_, oc), max(args.steps, 150))
        clocks = sampler.stop(t_a, time.perf_counter())
        redo = eng.last_dense_redo(B)
        flops = 2.0 * rows * dim * (((B + 31) // 32) * 32)
        print(json.dumps({"workload": f"cfg3: 10M x 1536 bf16, batch {B} shared segment (tcgen05 cta_group::2, N = batch rounded to 32)", "ms": ms,
                          "p50_ms": p50, "qps": B * 1e3 / ms, "tensor_TFLOPs": flops / (ms / 1e3) / 1e12,
                          "hbm_GBs": rows * dim * 2 / (ms / 1e3) / 1e9, "frac_hbm_measured": rows * dim * 2 / (ms / 1e3) / 1e9 / PEAK,
                          "uncertified_redone": redo, "launches": eng.last_search_launches, "clocks": clocks}), flush=True)
    eng.close(); del eng
    torch.cuda.empty_cache()


def fused_bench(args):
    rows, dim, k = args.rows, 1536, 10000
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=1, n_users=1)
    eng = Engine(device=0, max_batch=1, max_k=k)
    loader.collection_synthetic(spec, eng)
    qs = [torch.from_numpy(synth.query_vector(spec, i)[None, :]).to(DEV) for i in range(8)]
    oi = torch.empty((1, k), dtype=torch.int64, device=DEV); os_ = torch.empty((1, k), dtype=torch.float64, device=DEV)
    oc = torch.empty(1, dtype=torch.int32, device=DEV)
    fall = [make_filter(0, rows, synth.DATE_BASE, None)]
    mlo = (2022 - 1970) * 12
    ms_s, _ = timed(lambda i: eng.search_device(qs[i % 8], fall, k, oi, os_, oc), args.steps)
    out = {}
    def both(i):
        eng.search_device(qs[i % 8], fall, k, oi, os_, oc)
        out["r"] = eng.aggregate_hits(oi, oc, mlo, 24, 16)
    ms, p50 = timed(both, args.steps)
    sums, present, stats = out["r"]
    torch.cuda.synchronize()
    total = int(eng._keep[5][oi[0].long()].long().sum().item())      # ids = rows here: the cents of the returned points
    emit("fused: whale, k=10000 retrieval + (category, month) group-by of the returned points", ms, p50,
         rows * dim * 4 + rows * 4 + rows * 8, qps=1e3 / ms, search_only_ms=ms_s, groupby_ms=ms - ms_s,
         checksum_ok=bool(int(sums.sum().item()) == total and int(stats.sum().item()) == 0), groups=int(present.sum().item()))
    eng.close(); del eng
    torch.cuda.empty_cache()


def agg_bench(args):
    n, users = args.agg_rows, 100_000
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=users, seed=5, zipf_a=1.1)
    d = torch.empty(n, dtype=torch.int32, device=DEV)
    c = torch.empty(n, dtype=torch.uint8, device=DEV)
    v = torch.empty(n, dtype=torch.int32, device=DEV)
    synth_columns(0, n, spec.seed, 16, spec.date_span_seconds, dates=d, categories=c, cents=v)
    off = synth.segment_offsets(spec)
    eng = Engine(device=0, max_batch=1, max_k=1)
    # vectors are irrelevant for the aggregation but a shard needs them (row norms are scanned at load): 16 B/row of zeros
    vec = torch.zeros((n, 4), dtype=torch.float32, device=DEV)
    eng.load_shard(vec, dates=d, categories=c, cents=v, seg_offsets=torch.from_numpy(off).to(DEV))
    mlo = (2022 - 1970) * 12
    sums = torch.empty((users, 16, 24), dtype=torch.int64, device=DEV)
    pres = torch.empty((users, 16, 24), dtype=torch.uint8, device=DEV)
    drop = torch.empty(1, dtype=torch.int64, device=DEV)
    ref = int(v.to(torch.int64).sum().item())
    algo = n * 9 + users * 384 * 9          # 9 B per row (date 4, category 1, cents 4) + the dense output

    def run(label, kernel, **kw):
        if kernel:
            os.environ["FCB_AGG_KERNEL"] = kernel
        try:
            ms, p50 = timed(lambda i: eng.aggregate(0, users, mlo, 24, 16, sums, pres, drop), args.steps)
        finally:
            os.environ.pop("FCB_AGG_KERNEL", None)
        assert int(drop.item()) == 0
        emit(f"cfg5: {n} rows, 100k users x 16 categories x 24 months, {label}", ms, p50, algo, rows_per_s=n / (ms / 1e3),
             checksum_ok=(int(sums.sum().item()) == ref), groups=int(pres.sum().item()), **kw)

    run("dates in random order (group-key kernel: reads cents + 16-bit keys, 6 B/row)", "", bytes_read=n * 6 + users * 384 * 9,
        max_user_rows=int(np.diff(off).max()))
    if os.environ.get("FCB_BENCH_AGK_LAYOUTS"):
        os.environ["FCB_AGK_LAYOUT"] = "n"
        run("dates in random order (group-key kernel, NATURAL table layout)", "")
        os.environ.pop("FCB_AGK_LAYOUT")
        os.environ["FCB_AGK_CARVEOUT"] = "100"
        run("dates in random order (group-key kernel, maximum shared-memory carve-out)", "")
        os.environ.pop("FCB_AGK_CARVEOUT")
    run("dates in random order (date columns, 9 B/row: warp-table kernel, date form)", "date")
    run("dates in random order (date columns, 9 B/row: round 1's CTA-wide shared-atomic kernel + warp tables)", "cta")
    eng.close()
    # the same table as a time-ordered log (dates non-decreasing within each user)
    from financial_chatbot_llm_b200.engine import synth_dates_sorted
    offs_d = torch.from_numpy(off).to(DEV)
    synth_dates_sorted(d, offs_d, 0, spec.date_span_seconds)
    eng = Engine(device=0, max_batch=1, max_k=1)
    eng.load_shard(vec, dates=d, categories=c, cents=v, seg_offsets=offs_d)
    assert eng.dates_sorted
    run("time-ordered log (group-key kernel, 6 B/row)", "", bytes_read=n * 6 + users * 384 * 9)
    if os.environ.get("FCB_BENCH_AGK_LAYOUTS"):
        os.environ["FCB_AGK_LAYOUT"] = "n"
        run("time-ordered log (group-key kernel, NATURAL table layout)", "")
        os.environ.pop("FCB_AGK_LAYOUT")
    run("time-ordered log (date columns, 9 B/row: atomic-free runs)", "sorted")
    run("time-ordered log (date columns, 9 B/row: warp-table kernel, date form)", "date")
    eng.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--agg-rows", type=int, default=1_000_000_000)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    graft.build()
    if a.only in ("", "fused"):
        fused_bench(a)
    if a.only in ("", "agg"):
        agg_bench(a)
    if a.only in ("", "search"):
        search_bench(a)
