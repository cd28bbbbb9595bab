#!/usr/bin/env python
"""bench.py — filtered top-10 queries/s on BASELINE.json's configs[1]:
10M x 1536-d fp32 on one B200, batch 1, user_id + date-range filter, top-10.

  python bench.py [--gpus N] [--steps K] [--warmup W]          this framework (CUDA, C-ABI)
  python bench.py --impl reference [...]                       the reference path's CPU arithmetic
                                                               (oracle port of qdrant local mode)
A "step" is one pass of the hot path over one batch of synthetic queries:
  N=1  one query against the whale tenant's 10M-row segment (61.44 GB streamed from HBM).
  N>1  weak scaling by point range: rank r holds rows [r*10M, (r+1)*10M) = tenant r's segment;
       a step is a batch of N queries (one per tenant); every rank scans its shard for the
       whole batch, then ONE all_gather of the per-shard top-k candidates and a k-way merge.
value = queries/s with the queries resident in HBM (CUDA events, max over ranks).
e2e   = queries/s through B200QdrantClient.query_points with a host list[float] in and scored
        points (payloads included) out: host->device and device->host copies inside the timed region.
Inputs (61 GB per GPU) are ~500x the 126 MB L2, so no L2 flush is needed between steps.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

METRIC = "filtered_top10_queries_per_sec"
UNIT = "queries/s"
ROWS_PER_GPU = 10_000_000
DIM = 1536
K = 10
SEED = 20240607
# dram__bytes_read.sum + dram__bytes_write.sum per launch of scan_topk_kernel<float,2,6> on this workload, from the
# `ncu --set full` capture committed as profiles/r01_final_scan_topk_f32_2x6_full_raw.csv (61.5206 GB + 4.27 MB)
NCU_TRAFFIC_BYTES = 61_524_875_544   # 61.520607e9 + 4.268544e6


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (profiling recipe's line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, pw, reasons = [], [], [], set()
        for t, line in self.rows:
            if t0 is not None and not (t0 <= t <= t1 + 0.2):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            try:
                pw.append(float(f[3]))
            except ValueError:
                pass
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_median": statistics.median(pw) if pw else None}


# ------------------------------------------------------------------------------------------ reference arm
def host_sample(spec, rows: int, threads: int):
    """Rows [0, rows) of the SAME generator, on the host (thread pool over chunks)."""
    from concurrent.futures import ThreadPoolExecutor
    from financial_chatbot_llm_b200 import synth
    out = np.empty((rows, spec.dim), dtype=np.float32)
    chunk = 4096

    def fill(lo):
        hi = min(rows, lo + chunk)
        out[lo:hi] = synth.vectors_f32(spec, lo, hi)
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(fill, range(0, rows, chunk)))
    return out, synth.dates(spec, 0, rows)


def cpu_reference_qps(steps: int, warmup: int, sample_rows: int = 100_000, n_gpus: int = 1):
    """The reference path's arithmetic (qdrant-client local mode, restated in oracle/local_mode.py:
    per-query float32 re-normalisation of the stored matrix, mixed f32 x f64 dot over ALL rows,
    payload mask, descending order) timed on the host cores with numpy's default BLAS threading.
    Bounded sample: `sample_rows` rows of the same corpus; the scan is O(N), so queries/s for the
    10M-row workload = measured queries/s x sample_rows / 10M (stated in `sample`)."""
    from financial_chatbot_llm_b200 import synth
    from oracle import local_mode
    cores = os.cpu_count() or 1
    spec = synth.SynthSpec(n_rows=ROWS_PER_GPU, dim=DIM, seed=SEED, n_users=1)
    vec, dates = host_sample(spec, sample_rows, min(cores, 32))
    ids = np.arange(sample_rows, dtype=np.uint64)
    offsets = np.array([0, sample_rows], dtype=np.int64)
    t_lo = synth.DATE_BASE
    times = []
    for i in range(warmup + steps):
        q = synth.query_vector(spec, i).tolist()                 # list[float], as embed_query returns
        t = time.perf_counter()
        mask = local_mode.column_mask(sample_rows, offsets, 0, dates, t_lo, None, None, None)
        local_mode.search_columns(vec, ids, mask, q, K, inplace_renorm=True)
        dt = time.perf_counter() - t
        if i >= warmup:
            times.append(dt)
    qps_sample = len(times) / sum(times)
    # local mode scores EVERY stored row whatever the filter, so at N shards (N x 10M rows) a query costs N x more
    scale = sample_rows / (ROWS_PER_GPU * n_gpus)
    try:
        from threadpoolctl import threadpool_info
        blas_threads = max([p.get("num_threads", 1) for p in threadpool_info()] or [1])
    except Exception:
        blas_threads = None
    return {
        "value": qps_sample * scale, "unit": UNIT, "cores": cores, "kind": "port",
        "blas_threads": blas_threads,
        "sample": (f"{sample_rows} of {n_gpus}x10M rows x {DIM} f32, {len(times)} single queries, whale tenant + date>=t0 "
                   f"(100% pass), vectorised payload mask; measured {qps_sample:.3f} q/s on the sample, scaled by "
                   f"{scale:g} (scan is O(N)); numpy {np.__version__}"),
        "ms_per_query_sample": 1e3 * statistics.median(times),
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_reference_qps(args.steps, args.warmup, n_gpus=args.gpus)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / cb["value"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference arithmetic = qdrant-client 1.14.2 local mode; the package is not installable here, "
                "so this is the oracle port (oracle/local_mode.py) on the host cores",
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus: int):
    return {"workload": "configs[1]: 10M x 1536 fp32 per GPU, batch 1 per tenant, user_id + date-range filter, top-10",
            "rows_per_gpu": ROWS_PER_GPU, "dim": DIM, "k": K, "storage": "fp32",
            "filter": "metadata.user_id == whale tenant AND metadata.date >= t0 (all of the tenant's rows pass)",
            "queries_per_step": n_gpus, "sharding": "point range, 1 tenant segment per GPU" if n_gpus > 1 else "none",
            "l2": "inputs 61 GB per GPU >> 126 MB L2: no flush between steps", "seed": SEED}


# ------------------------------------------------------------------------------------------ B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as graft
    from financial_chatbot_llm_b200 import _lib, loader, synth
    from financial_chatbot_llm_b200 import qdrant_models as models
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.engine import Engine, make_filter

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ["NCCL_DEBUG"] = os.environ.get("FCB_NCCL_DEBUG", "WARN")   # keep stdout to the ONE JSON line
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        graft.build()
    if world > 1:
        dist.barrier()
    _lib.load()

    rows = args.rows
    n_total = rows * world
    spec = synth.SynthSpec(n_rows=n_total, dim=DIM, seed=SEED, n_users=world, zipf_a=0.0)
    eng = Engine(device=local_rank, max_batch=max(world, 1), max_k=16)
    t_gen = time.perf_counter()
    col = loader.collection_synthetic(spec, eng, row_lo=rank * rows, row_hi=(rank + 1) * rows, with_cents=False)
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t_gen
    offsets = col.offsets
    B = world
    t_lo = synth.DATE_BASE                                    # every date >= t0: predicate evaluated, all pass
    filters = [make_filter(int(offsets[u]), int(offsets[u + 1]), t_lo, None) for u in range(B)]

    n_q = 16
    q_host = [np.stack([synth.query_vector(spec, s * B + b) for b in range(B)]) for s in range(n_q)]
    q_dev = [torch.from_numpy(q).to(dev) for q in q_host]
    from financial_chatbot_llm_b200 import sharded
    searcher = sharded.ShardedSearcher(col, world, dev)
    merger = searcher.merger(B, K)
    out_ids, out_sc, out_cnt = merger.local_views          # fcb_search writes straight into the packed buffer

    def step(i):
        eng.search_device(q_dev[i % n_q], filters, K, out_ids, out_sc, out_cnt)
        if world > 1:
            merger.merge()                                  # one all_gather_into_tensor + fcb_merge_topk

    launches_per_step = 1 + (1 if world > 1 else 0)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    time.sleep(0.3 if rank == 0 else 0.0)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    t0 = time.perf_counter()
    evs[0].record()
    for i in range(args.steps):
        step(i)
        evs[i + 1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t1 = time.perf_counter()
    total_ms = evs[0].elapsed_time(evs[-1])
    per_step = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps))
    tm = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    total_ms = float(tm.item())
    clocks = sampler.stop(t0, t1) if sampler else None

    # ---- correctness spot check of the timed configuration (size-independent property + prefix oracle)
    final_ids = (merger.ids if world > 1 else out_ids).cpu().numpy().view(np.uint64)
    final_sc = (merger.scores if world > 1 else out_sc).cpu().numpy()
    assert np.all(np.diff(final_sc, axis=1) <= 0), "scores must be non-increasing"
    for b in range(B):                                      # every hit belongs to the queried tenant's segment
        assert np.all((final_ids[b] >= offsets[b]) & (final_ids[b] < offsets[b + 1]))

    # ---- e2e: through the public API with HOST buffers in and out.
    #   N=1: B200QdrantClient.query_points (list[float] in, ScoredPoint + payload out).
    #   N>1: ShardedSearcher.search (numpy batch in, numpy results out; collective on every rank).
    t_lo_filter = t_lo
    e2e_steps = args.steps
    if world == 1:
        client = B200QdrantClient(url="unused")
        client.attach_collection(col)
        flt = models.Filter(must=[
            models.FieldCondition(key="metadata.user_id", match=models.MatchValue(value=col.users[0])),
            models.FieldCondition(key="metadata.date", range=models.Range(gte=t_lo_filter))])
        q_lists = [q_host[s][0].tolist() for s in range(n_q)]
        for i in range(3):
            client.query_points("transactions", query=q_lists[i % n_q], limit=K, query_filter=flt)
        te = time.perf_counter()
        for i in range(e2e_steps):
            pts = client.query_points("transactions", query=q_lists[i % n_q], limit=K, query_filter=flt).points
        te = time.perf_counter() - te
        assert len(pts) == K
        e2e_api = "B200QdrantClient.query_points (list[float] in, ScoredPoint+payload out)"
    else:
        for i in range(3):
            searcher.search(q_host[i % n_q], filters, K)
        dist.barrier()
        te = time.perf_counter()
        for i in range(e2e_steps):
            r_ids, r_sc, r_cnt = searcher.search(q_host[i % n_q], filters, K)
        te = time.perf_counter() - te
        assert int(r_cnt.min()) == K
        e2e_api = "ShardedSearcher.search (numpy [B,dim] in, numpy ids/scores/counts out, collective)"
    tt = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    te = float(tt.item())
    kernel_ms_e2e = eng.last_search_ms

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ms_per_step = total_ms / args.steps
    qps = B * args.steps / (total_ms / 1e3)
    peak, peak_src = measured_peaks()
    seg_rows = rows
    alg_bytes = seg_rows * DIM * 4 + seg_rows * 4              # vectors of passing rows + date predicate column
    kern_ms = ms_per_step                                      # N>1: scan + all-gather + merge (upper bound on the scan)
    achieved = alg_bytes / (kern_ms / 1e3) / 1e9
    line = {
        "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(world),
        "p50_ms": per_step[len(per_step) // 2], "p99_ms": per_step[min(len(per_step) - 1, int(len(per_step) * 0.99))],
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": NCU_TRAFFIC_BYTES if rows == ROWS_PER_GPU else None, "kernel": "scan_topk_kernel<float,%d,%d>" % (eng.scan_config if eng.scan_config != (0, 0) else (2, 6)), "algorithmic_bytes_per_launch": alg_bytes,
                     "kernel_ms": kern_ms, "peak_source": peak_src, "frac_of_nominal_8TBs": achieved / 8000.0},
        "e2e": {"value": B * e2e_steps / te, "unit": UNIT, "h2d_bytes_per_step": B * DIM * 8,
                "d2h_bytes_per_step": B * (K * 16 + 4), "api": e2e_api,
                "ms_per_query": 1e3 * te / e2e_steps, "kernel_ms": kernel_ms_e2e},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
        "corpus_gen_s": t_gen,
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_reference_qps(steps=8, warmup=1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--rows", type=int, default=ROWS_PER_GPU, help="rows per GPU (default: the BASELINE config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
