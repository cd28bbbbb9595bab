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
        points (payloads included) out: host->device and device->host copies inside the timed region;
        p50 / p99 of the per-call wall-clock latency are reported with it.
Inputs (61 GB per GPU) are ~500x the 126 MB L2, so no L2 flush is needed between steps.

After the headline (same JSON line, the headline keys untouched):
  N=1  "extra": BASELINE configs[2] (10M x 1536 bf16, batch 256: tcgen05 path, tensor roofline) and configs[4]
       (1B-row integer-cents group-by by user x category x month, any row order: HBM roofline), each with its own
       clock record and a parity check against the oracle on host-regenerated rows.
  N>1  "cfg4": BASELINE configs[3] (100M x 1024 bf16 sharded by point range over the N GPUs, 100k tenants, batches of 64
       distinct tenants; and one whale query striped over all shards), with a parity flag.
"parity_spot_check": the timed configuration's answer re-scored from host-regenerated rows by the C oracle (bit for
bit) and checked against a host sample of every shard (no sampled row beats the k-th).
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
REF_SAMPLE_ROWS = 1_000_000          # rows the CPU arm really scores per query (scale to the 10M workload: 0.1)


def measured_peaks():
    """(hbm GB/s, bf16 sustained TFLOP/s, source)"""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            return float(j["hbm_gbs"]), float(j.get("bf16_tflops_sustained", 1370.7)), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, 1400.0, "fallback (B200_PROFILING.md: 6.65 TB/s copy, ~1.4 PFLOP/s sustained bf16)"


def ncu_traffic(kernel_prefix: str):
    """dram bytes per launch of the dominant kernel from a committed `ncu --set full` capture, if one is on record
    (profiles/ncu_traffic.json, written by scripts/ncu_summary.py from the capture named in it); else None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        j = json.load(open(p))
        for name, rec in j.items():
            if name.startswith(kernel_prefix):
                return rec
    except Exception:
        pass
    return None


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


def pct(sorted_vals, q):
    return sorted_vals[min(len(sorted_vals) - 1, int(len(sorted_vals) * q))]


# ------------------------------------------------------------------------------------------ reference arm
def host_sample(spec, row_lo: int, rows: int, threads: int, bf16: bool = False):
    """Rows [row_lo, row_lo + rows) of the SAME generator, on the host (thread pool over chunks)."""
    from concurrent.futures import ThreadPoolExecutor
    from financial_chatbot_llm_b200 import synth
    out = np.empty((rows, spec.dim), dtype=np.uint16 if bf16 else np.float32)
    chunk = 4096

    def fill(lo):
        hi = min(rows, lo + chunk)
        if bf16:
            out[lo:hi] = synth.vectors_bf16_bits(spec, row_lo + lo, row_lo + hi).reshape(hi - lo, spec.dim)
        else:
            out[lo:hi] = synth.vectors_f32(spec, row_lo + lo, row_lo + hi)
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(fill, range(0, rows, chunk)))
    return out


def cpu_reference_qps(steps: int, warmup: int, sample_rows: int = REF_SAMPLE_ROWS, n_gpus: int = 1, vec=None):
    """The reference path's arithmetic (qdrant-client local mode, restated in oracle/local_mode.py:
    per-query float32 re-normalisation of the stored matrix, mixed f32 x f64 dot over ALL rows,
    payload mask, descending order) timed on the host cores with every BLAS thread the box has (torchrun exports
    OMP_NUM_THREADS=1 to its ranks; the thread pool is raised back to the core count here).
    Bounded sample: `sample_rows` rows of the same corpus are really scored per query; the scan is O(N), so
    queries/s for the 10M-row workload = measured queries/s x sample_rows / 10M (both numbers are reported)."""
    from financial_chatbot_llm_b200 import synth
    from oracle import local_mode
    cores = os.cpu_count() or 1
    limiter = None
    try:
        from threadpoolctl import threadpool_limits
        limiter = threadpool_limits(limits=cores)
    except Exception:
        pass
    spec = synth.SynthSpec(n_rows=ROWS_PER_GPU, dim=DIM, seed=SEED, n_users=1)
    if vec is None:                                             # (the B200 arm hands over rows of the same generator, made on the device)
        vec = host_sample(spec, 0, sample_rows, min(cores, 32))
    sample_rows = vec.shape[0]
    dates = synth.dates(spec, 0, sample_rows)
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
    del limiter
    return {
        "value": qps_sample * scale, "unit": UNIT, "cores": cores, "kind": "port",
        "blas_threads": blas_threads, "measured_qps_on_sample": qps_sample, "sample_rows": sample_rows, "scale_to_workload": scale,
        "sample": (f"{sample_rows} of {n_gpus}x10M rows x {DIM} f32 really scored per query, {len(times)} single queries, whale "
                   f"tenant + date>=t0 (100% pass), vectorised payload mask; measured {qps_sample:.4f} q/s on the sample, "
                   f"x {scale:g} for the full row count (the scan is O(N)); numpy {np.__version__}"),
        "ms_per_query_sample": 1e3 * statistics.median(times),
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded: 1M rows take ~5 s per query on 16 cores; cap the number of timed queries so the run ends within minutes
    steps = max(1, min(args.steps, 12))
    warmup = max(1, min(args.warmup, 2))
    cb = cpu_reference_qps(steps, warmup, sample_rows=args.ref_rows, n_gpus=args.gpus)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        # a step of this arm = one query over the bounded sample (what was really timed); `value` scales it to the workload
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 / cb["measured_qps_on_sample"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference arithmetic = qdrant-client 1.14.2 local mode; the package is not installable here, "
                "so this is the oracle port (oracle/local_mode.py) on the host cores. steps/warmup are capped (12/2): "
                "a query over the 1M-row sample takes seconds",
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus: int):
    return {"workload": "configs[1]: 10M x 1536 fp32 per GPU, batch 1 per tenant, user_id + date-range filter, top-10",
            "rows_per_gpu": ROWS_PER_GPU, "dim": DIM, "k": K, "storage": "fp32",
            "filter": "metadata.user_id == whale tenant AND metadata.date >= t0 (all of the tenant's rows pass)",
            "queries_per_step": n_gpus, "sharding": "point range, 1 tenant segment per GPU" if n_gpus > 1 else "none",
            "l2": "inputs 61 GB per GPU >> 126 MB L2: no flush between steps", "seed": SEED}


# ------------------------------------------------------------------------------------------ timing helper
def timed_steps(step, steps, warmup, dev, world, sample_gpu=None):
    """W warm-ups, barrier + sync, K steps between CUDA events on the current stream, barrier + sync; returns
    (ms per step as max over ranks, sorted per-step ms of this rank, clocks or None)."""
    import torch
    import torch.distributed as dist
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(sample_gpu).start() if sample_gpu is not None else None
    if sampler:
        time.sleep(0.3)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    t0 = time.perf_counter()
    evs[0].record()
    for i in range(steps):
        step(i)
        evs[i + 1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t1 = time.perf_counter()
    tm = torch.tensor([evs[0].elapsed_time(evs[-1])], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
    return float(tm.item()) / steps, per, (sampler.stop(t0, t1) if sampler else None)


# ------------------------------------------------------------------------------------------ extras (N = 1)
def extra_cfg3(dev, steps):
    """BASELINE configs[2]: 10M x 1536 bf16, 256 concurrent queries on one shared segment (tcgen05 / TMEM path)."""
    import torch
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine, make_filter
    from oracle import canon
    rows, B = ROWS_PER_GPU, 256
    spec = synth.SynthSpec(n_rows=rows, dim=DIM, seed=3, n_users=1)
    eng = Engine(device=dev.index, max_batch=B, max_k=16)
    loader.collection_synthetic(spec, eng, bf16=True, with_cents=False)
    qs = np.stack([synth.query_vector(spec, 500 + i) for i in range(B)])
    q = torch.from_numpy(qs).to(dev)
    f = [make_filter(0, rows)] * B
    oi = torch.empty((B, K), dtype=torch.int64, device=dev)
    os_ = torch.empty((B, K), dtype=torch.float64, device=dev)
    oc = torch.empty(B, dtype=torch.int32, device=dev)
    ms, per, clocks = timed_steps(lambda i: eng.search_device(q, f, K, oi, os_, oc), steps, 3, dev, 1, sample_gpu=dev.index)
    redo, reruns = eng.last_dense_redo(B), eng.exact_reruns
    ids, sc = oi.cpu().numpy().view(np.uint64), os_.cpu().numpy()
    ok = bool((oc.cpu().numpy() == K).all())
    for b in range(0, B, 32):                                       # returned rows re-scored by the C oracle from host-regenerated rows
        for j in (0, K - 1):
            r = int(ids[b, j])
            ok = ok and canon.scores(synth.vectors_bf16_bits(spec, r, r + 1).reshape(1, DIM), qs[b])[0] == sc[b, j]
        ok = ok and bool(np.all(np.diff(sc[b]) <= 0))
    blk = host_sample(spec, 6_000_000, 50_000, min(os.cpu_count() or 1, 32), bf16=True)
    for b in (0, 128):                                              # no row of a host sample beats the k-th result
        better = np.flatnonzero(canon.scores(blk, qs[b]) > sc[b, K - 1]) + 6_000_000
        ok = ok and set(better.tolist()) <= set(ids[b].astype(np.int64).tolist())
    eng.close()
    del eng, q, oi, os_, oc
    torch.cuda.empty_cache()
    hbm_peak, tf_peak, src = measured_peaks()
    flops = 2.0 * rows * DIM * B
    tf = flops / (ms / 1e3) / 1e12
    return {"workload": "configs[2]: 10M x 1536 bf16, batch 256 concurrent queries on one shared segment, top-10, float64 rescore",
            "ms": ms, "p50_ms": per[len(per) // 2], "queries_per_s": B * 1e3 / ms, "steps": steps, "gpu_launches_per_step": 5,
            "kernel": "dense_topk2_kernel (tcgen05.mma cta_group::2, TMA, TMEM) + dense_finalize_kernel",
            "roofline": {"bound": "tensor", "achieved": tf, "peak": tf_peak, "unit": "TFLOP/s", "frac": tf / tf_peak,
                         "flops_per_launch": flops, "peak_source": src + " bf16_tflops_sustained", "traffic": None,
                         "hbm_GBs": rows * DIM * 2 / (ms / 1e3) / 1e9, "hbm_frac": rows * DIM * 2 / (ms / 1e3) / 1e9 / hbm_peak},
            "uncertified_rerun_by_scan": redo, "answered_by_float64_path": reruns, "clocks": clocks, "parity": ok,
            "parity_how": "16 returned rows re-scored bit for bit by the C oracle from host-regenerated bf16 rows; no row of a 50k-row host "
                          "sample beats the 10th result of 2 queries; full property test: tests/test_gpu_fullsize_dense.py"}


def extra_cfg5(dev, steps):
    """BASELINE configs[4]: 1B-row integer-cents group-by by (user, category, month), rows in any order."""
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import aggregate_columns, build_group_keys, synth_columns
    from oracle import groupby as ogroupby
    n, users = 1_000_000_000, 100_000
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=users, seed=5, zipf_a=1.1)
    d = torch.empty(n, dtype=torch.int32, device=dev)
    c = torch.empty(n, dtype=torch.uint8, device=dev)
    v = torch.empty(n, dtype=torch.int32, device=dev)
    synth_columns(0, n, spec.seed, 16, spec.date_span_seconds, dates=d, categories=c, cents=v)
    off = synth.segment_offsets(spec)
    offs_d = torch.from_numpy(off).to(dev)
    mlo = (2022 - 1970) * 12
    sums = torch.empty((users, 16, 24), dtype=torch.int64, device=dev)
    pres = torch.empty((users, 16, 24), dtype=torch.uint8, device=dev)
    drop = torch.empty(1, dtype=torch.int64, device=dev)
    # The line's own number is the RAW-COLUMN run — dates, categories, cents as the reference holds them, 9 B/row.
    # A resident table can in addition carry group keys ((category, month) of every row in 16 bits, resolved once like the row
    # norms of the search shard): the group-by then reads 6 B per row; that run is reported beside it under "with_group_keys".
    torch.cuda.synchronize(dev)
    t_k = time.perf_counter()
    keys, key_base = build_group_keys(dev.index, d, c, v)
    torch.cuda.synchronize(dev)
    keys_ms = 1e3 * (time.perf_counter() - t_k)                  # one-time cost per resident table (two passes over the columns)
    run_keyed = lambda i: aggregate_columns(dev.index, d, c, v, offs_d, 0, users, mlo, 24, 16, sums, pres, drop, keys=keys, key_base=key_base)
    ms_keyed, _, _ = timed_steps(run_keyed, max(3, steps // 2), 3, dev, 1)
    sums_dates, pres_dates = sums.clone(), pres.clone()              # (the keyed result, compared with the raw-column one below)
    run = lambda i: aggregate_columns(dev.index, d, c, v, offs_d, 0, users, mlo, 24, 16, sums, pres, drop)
    ms, per, clocks = timed_steps(run, steps, 3, dev, 1, sample_gpu=dev.index)
    same = bool(torch.equal(sums, sums_dates)) and bool(torch.equal(pres, pres_dates))
    del sums_dates, pres_dates
    ok = same and int(drop.item()) == 0 and int(sums.sum().item()) == int(v.to(torch.int64).sum().item())
    sizes = np.diff(off)
    for u in (int(np.argmin(np.abs(sizes - 200_000))), 4321, users - 1):     # three users against pandas on host-regenerated columns
        a, b = int(off[u]), int(off[u + 1])
        want = ogroupby.groupby_user_category_month(np.full(b - a, u), synth.categories(spec, a, b), synth.dates(spec, a, b),
                                                    synth.cents(spec, a, b))
        got_s, got_p = sums[u].cpu().numpy(), pres[u].cpu().numpy().astype(bool)
        cells = {(int(cc), int(mm) + mlo): int(got_s[cc, mm]) for cc, mm in zip(*np.nonzero(got_p))}
        ok = ok and cells == {(int(k[1]), int(k[2])): int(x) for k, x in want.items()}
    del d, c, v, sums, pres, keys
    torch.cuda.empty_cache()
    hbm_peak, _, src = measured_peaks()
    alg = n * 9 + users * 384 * 9            # date 4 + category 1 + cents 4 bytes per row, + the dense output
    read_keyed = n * 6 + users * 384 * 9     # what the keyed form really moves: cents 4 + key 2 bytes per row
    gbs = alg / (ms / 1e3) / 1e9
    return {"workload": "configs[4]: 1B rows, int32 cents summed by user x category x month (100k Zipf users x 16 x 24), rows in any order",
            "ms": ms, "p50_ms": per[len(per) // 2], "rows_per_s": n / (ms / 1e3), "steps": steps,
            "kernel": "aggregate_warp_kernel, date form: dates / categories / cents as the reference holds them (9 B/row)",
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
                         "algorithmic_bytes_per_launch": alg, "peak_source": src + " hbm_gbs",
                         "traffic": (ncu_traffic("aggregate_warp_kernel<date>") or {}).get("dram_bytes_per_launch")},
            "with_group_keys": {"ms": ms_keyed, "kernel": "aggregate_warp_kernel, keyed form: cents + 16-bit group keys (6 B/row)",
                                "group_keys_build_ms_once": keys_ms, "bytes_read_per_launch": read_keyed,
                                "read_GBs": read_keyed / (ms_keyed / 1e3) / 1e9, "read_frac": read_keyed / (ms_keyed / 1e3) / 1e9 / hbm_peak,
                                "frac_in_9B_per_row_terms": alg / (ms_keyed / 1e3) / 1e9 / hbm_peak,
                                "traffic": (ncu_traffic("aggregate_warp_kernel<keyed>") or {}).get("dram_bytes_per_launch")},
            "clocks": clocks, "parity": bool(ok),
            "parity_how": "keyed and date-column kernels give the identical dense output, no row dropped, total of all cells == total "
                          "of the cents column, and every cell of three users (up to 200k rows) equal to pandas groupby().sum() on "
                          "host-regenerated columns; bit-exact tests: tests/test_gpu_aggregate.py"}


# ------------------------------------------------------------------------------------------ cfg4 (N > 1)
def bench_cfg4(dev, world, rank, steps):
    import bench_cfg4
    return bench_cfg4.run(dev, world, rank, steps, timed_steps)


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
        # NCCL's INFO lines (communicator sizes, NVLS) stay observable but go to stderr: stdout carries the ONE JSON line
        os.environ.setdefault("NCCL_DEBUG", "INFO")
        os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
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

    step(0)
    torch.cuda.synchronize()
    launches_per_step = eng.last_search_launches + (1 if world > 1 else 0)   # scan + flag-gated float64 path (+ merge)

    ms_per_step, per_step, clocks = timed_steps(step, args.steps, args.warmup, dev, world, sample_gpu=local_rank if rank == 0 else None)
    total_ms = ms_per_step * args.steps
    last_q = (args.steps - 1) % n_q

    # ---- parity spot check of the timed configuration: the answer of the LAST timed step against the C oracle on
    #      host-regenerated rows (bit for bit), order and segment membership, and a 100k-row host sample of THIS shard
    from oracle import canon
    final_ids = (merger.ids if world > 1 else out_ids).cpu().numpy().view(np.uint64)
    final_sc = (merger.scores if world > 1 else out_sc).cpu().numpy()
    parity = bool(np.all(np.diff(final_sc, axis=1) <= 0))
    for b in range(B):
        parity = parity and bool(np.all((final_ids[b] >= offsets[b]) & (final_ids[b] < offsets[b + 1])))
        for j in (0, K // 2, K - 1):
            r = int(final_ids[b, j])
            parity = parity and canon.scores(synth.vectors_f32(spec, r, r + 1), q_host[last_q][b])[0] == final_sc[b, j]
    lo_s = rank * rows + (rows // 3)
    blk = host_sample(spec, lo_s, min(100_000, rows - rows // 3), min(os.cpu_count() or 1, 32 // max(1, min(world, 8)) or 1))
    s_blk = canon.scores(blk, q_host[last_q][rank])          # this rank's tenant = this rank's shard
    better = np.flatnonzero(s_blk > final_sc[rank, K - 1]) + lo_s
    parity = parity and set(better.tolist()) <= set(final_ids[rank].astype(np.int64).tolist())
    parity = parity and eng.exact_reruns == 0
    if world > 1:
        pt = torch.tensor([1 if parity else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(pt, op=dist.ReduceOp.MIN)
        parity = bool(pt.item())

    # ---- e2e: through the public API with HOST buffers in and out.
    #   N=1: B200QdrantClient.query_points (list[float] in, ScoredPoint + payload out).
    #   N>1: ShardedSearcher.search (numpy batch in, numpy results out; collective on every rank).
    e2e_steps = args.steps
    lat = []
    if world == 1:
        client = B200QdrantClient(url="unused")
        client.attach_collection(col)
        flt = models.Filter(must=[
            models.FieldCondition(key="metadata.user_id", match=models.MatchValue(value=col.users[0])),
            models.FieldCondition(key="metadata.date", range=models.Range(gte=t_lo))])
        q_lists = [q_host[s][0].tolist() for s in range(n_q)]
        for i in range(3):
            client.query_points("transactions", query=q_lists[i % n_q], limit=K, query_filter=flt)
        te = time.perf_counter()
        for i in range(e2e_steps):
            t1 = time.perf_counter()
            pts = client.query_points("transactions", query=q_lists[i % n_q], limit=K, query_filter=flt).points
            lat.append(time.perf_counter() - t1)
        te = time.perf_counter() - te
        assert len(pts) == K
        e2e_api = "B200QdrantClient.query_points (list[float] in, ScoredPoint+payload out)"
    else:
        for i in range(3):
            searcher.search(q_host[i % n_q], filters, K)
        dist.barrier()
        te = time.perf_counter()
        for i in range(e2e_steps):
            t1 = time.perf_counter()
            r_ids, r_sc, r_cnt = searcher.search(q_host[i % n_q], filters, K)
            lat.append(time.perf_counter() - t1)
        te = time.perf_counter() - te
        assert int(r_cnt.min()) == K
        e2e_api = "ShardedSearcher.search (numpy [B,dim] in, numpy ids/scores/counts out, collective)"
    tt = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    te = float(tt.item())
    kernel_ms_e2e = eng.last_search_ms
    lat.sort()
    scan_cfg = eng.scan_config

    # ---- the scan alone (N > 1: the step above also holds the all-gather and the merge): CUDA events around fcb_search only
    scan_ms = ms_per_step
    if world > 1:
        scan_ms, _, _ = timed_steps(lambda i: eng.search_device(q_dev[i % n_q], filters, K, out_ids, out_sc, out_cnt),
                                    max(5, args.steps // 2), 2, dev, world)

    # ---- free the headline corpus, then the other BASELINE configs
    eng.close()
    del eng, col, searcher, merger, out_ids, out_sc, out_cnt, q_dev
    torch.cuda.empty_cache()
    extra, cfg4_res = None, None
    if not args.headline_only:
        if world == 1:
            extra = []
            for fn in (extra_cfg3, extra_cfg5):
                try:
                    extra.append(fn(dev, 20 if fn is extra_cfg3 else 10))
                except Exception as e:                           # the headline line must survive a failing extra
                    extra.append({"workload": fn.__name__, "error": f"{type(e).__name__}: {e}"})
                torch.cuda.empty_cache()
        else:
            try:
                cfg4_res = bench_cfg4(dev, world, rank, max(20, args.steps))
            except Exception as e:
                cfg4_res = {"error": f"{type(e).__name__}: {e}"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    qps = B * args.steps / (total_ms / 1e3)
    hbm_peak, _, peak_src = measured_peaks()
    alg_bytes = rows * DIM * 4 + rows * 4                       # vectors of passing rows + date predicate column
    achieved = alg_bytes / (scan_ms / 1e3) / 1e9
    kname = "scan_topk_kernel<float,%d,%d>" % (scan_cfg if scan_cfg != (0, 0) else (2, 6))
    # (every loop shape reads every row once: the capture of one shape stands for the others the load-time tuner may pick)
    traffic = (ncu_traffic(kname) or ncu_traffic("scan_topk_kernel<float")) if rows == ROWS_PER_GPU else None
    line = {
        "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(world),
        "p50_ms": per_step[len(per_step) // 2], "p99_ms": pct(per_step, 0.99),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                     "traffic": traffic["dram_bytes_per_launch"] if traffic else None,
                     "traffic_source": traffic["capture"] if traffic else None,
                     "kernel": kname, "algorithmic_bytes_per_launch": alg_bytes,
                     "kernel_ms": scan_ms, "kernel_ms_how": "CUDA events around the scan launches only" if world > 1 else "= ms_per_step (one launch per step)",
                     "peak_source": peak_src + " hbm_gbs", "frac_of_nominal_8TBs": achieved / 8000.0},
        "e2e": {"value": B * e2e_steps / te, "unit": UNIT, "h2d_bytes_per_step": B * DIM * 8,
                "d2h_bytes_per_step": B * (K * 16 + 4), "api": e2e_api,
                "ms_per_query": 1e3 * te / e2e_steps, "p50_ms": 1e3 * lat[len(lat) // 2], "p99_ms": 1e3 * pct(lat, 0.99),
                "kernel_ms": kernel_ms_e2e},
        "gpu_launches": launches_per_step * args.steps,
        "parity_spot_check": parity,
        "clocks": clocks,
        "corpus_gen_s": t_gen,
    }
    if extra is not None:
        line["extra"] = extra
    if cfg4_res is not None:
        line["cfg4"] = cfg4_res
    if world == 1 and not args.no_cpu_baseline:
        from financial_chatbot_llm_b200.engine import synth_vectors
        n_s = min(args.ref_rows, rows)
        sample = torch.empty((n_s, DIM), dtype=torch.float32, device=dev)   # rows [0, n_s) of the same corpus, bit-identical to the host generator
        synth_vectors(sample, 0, SEED, spec.n_clusters)
        vec = sample.cpu().numpy()
        del sample
        line["cpu_baseline"] = cpu_reference_qps(steps=3, warmup=1, vec=vec)
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
    ap.add_argument("--headline-only", action="store_true", help="skip the extra BASELINE configs (extra / cfg4)")
    ap.add_argument("--ref-rows", type=int, default=REF_SAMPLE_ROWS, help="rows the CPU arm really scores per query")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
