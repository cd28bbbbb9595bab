#!/usr/bin/env python
"""BASELINE config 4: 100M x 1024-d bf16 sharded by point range across G GPUs of one box, batch 64,
100k-tenant user_id filter, one all-gather + k-way merge per batch. Launch with torchrun:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 scripts/bench_cfg4.py
Prints one JSON line per sub-config on rank 0:
  tenants   B = 64 distinct random tenants per step (~1000 rows each): latency-bound
  whale     one tenant striped over all shards (N/G rows per GPU): bandwidth scaling
Timing: CUDA events, max over ranks; parity spot-check against the host generator + C oracle."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=100_000_000)
ap.add_argument("--dim", type=int, default=1024)
ap.add_argument("--tenants", type=int, default=100_000)
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--check", type=int, default=4)
args = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
import __graft_entry__ as graft
if rank == 0:
    graft.build()
if world > 1:
    dist.barrier()
from financial_chatbot_llm_b200 import loader, sharded, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter

K = 10
spec = synth.SynthSpec(n_rows=args.rows, dim=args.dim, seed=44, n_users=args.tenants, zipf_a=0.0)   # ~equal tenants
offsets = synth.segment_offsets(spec)
lo, hi = sharded.shard_ranges(args.rows, world, offsets)[rank]
eng = Engine(device=lr, max_batch=args.batch, max_k=16)
t0 = time.perf_counter()
col = loader.collection_synthetic(spec, eng, bf16=True, row_lo=lo, row_hi=hi, with_cents=False)
torch.cuda.synchronize()
gen_s = time.perf_counter() - t0
searcher = sharded.ShardedSearcher(col, world, dev)


def timed(step, steps, warmup=5):
    for i in range(warmup): step(i)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        step(i); evs[i + 1].record()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    tot = torch.tensor([evs[0].elapsed_time(evs[-1])], dtype=torch.float64, device=dev)
    if world > 1: dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
    return float(tot.item()) / steps, per[len(per) // 2]


# ---- tenants: B distinct random tenants per step
rng = np.random.default_rng(7)
B = args.batch
n_sets = 32
tsets = [rng.choice(args.tenants, size=B, replace=False) for _ in range(n_sets)]
fsets = [[make_filter(int(offsets[t]), int(offsets[t + 1])) for t in ts] for ts in tsets]
qsets = [torch.from_numpy(np.stack([synth.query_vector(spec, s * B + b) for b in range(B)])).to(dev) for s in range(4)]
m = searcher.merger(B, K)
oi, osc, oc = m.local_views

def step_tenants(i):
    eng.search_device(qsets[i % 4], fsets[i % n_sets], K, oi, osc, oc)
    m.merge()

ms, p50 = timed(step_tenants, args.steps)
seg_rows = float(np.mean([offsets[t + 1] - offsets[t] for t in tsets[0]]))
# parity spot-check: regenerate a few tenants on the host and compare with the C oracle
ok = True
if args.check:
    from oracle import canon
    step_tenants(0); torch.cuda.synchronize()
    ids = m.ids.cpu().numpy().view(np.uint64); sc = m.scores.cpu().numpy(); cnt = m.counts.cpu().numpy()
    for b in range(args.check):
        t = int(tsets[0][b]); a, e_ = int(offsets[t]), int(offsets[t + 1])
        bits = synth.vectors_bf16_bits(spec, a, e_).reshape(e_ - a, args.dim)
        q = synth.query_vector(spec, b)
        oi_, os_ = canon.search(bits, q, K, id_base=a)
        ok &= ids[b, :cnt[b]].tolist() == oi_.tolist() and sc[b, :cnt[b]].tolist() == os_.tolist()
if rank == 0:
    print(json.dumps({"workload": f"cfg4 tenants: {args.rows} x {args.dim} bf16 over {world} GPUs, batch {B} distinct tenants of {args.tenants}",
                      "n_gpus": world, "ms_per_batch": ms, "p50_ms": p50, "qps": B * 1e3 / ms, "mean_segment_rows": seg_rows,
                      "algorithmic_bytes_per_batch": seg_rows * B * args.dim * 2, "allgather_bytes_per_rank": sharded.packed_len(B, K) * 8,
                      "parity_spot_check": bool(ok), "rows_per_gpu": hi - lo, "corpus_gen_s": gen_s}), flush=True)

# ---- whale: one query whose segment is the whole corpus (striped over all shards)
qw = torch.from_numpy(synth.query_vector(spec, 999)[None, :]).to(dev)
m1 = searcher.merger(1, K)
wi, ws, wc = m1.local_views
fw = [make_filter(0, args.rows)]

def step_whale(i):
    eng.search_device(qw, fw, K, wi, ws, wc)
    m1.merge()

ms, p50 = timed(step_whale, max(20, args.steps // 10))
if rank == 0:
    gbs = args.rows * args.dim * 2 / (ms / 1e3) / 1e9
    print(json.dumps({"workload": f"cfg4 whale: one query over all {args.rows} rows striped over {world} GPUs", "n_gpus": world,
                      "ms_per_query": ms, "p50_ms": p50, "aggregate_GBs": gbs, "per_gpu_GBs": gbs / world}), flush=True)
if world > 1:
    dist.destroy_process_group()
