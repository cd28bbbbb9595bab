#!/usr/bin/env python
"""Small tenants (<= 1024 rows): fused scan (k = 10) against the direct float64 path (k = 17), batch 1 and 64."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter
DEV = torch.device("cuda", 0)
rows, dim = 10_000_000, 1536
spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=2, n_users=10000, zipf_a=1.1)
eng = Engine(device=0, max_batch=64, max_k=10000)
col = loader.collection_synthetic(spec, eng, with_cents=False)
off = col.offsets
rng = np.random.default_rng(0)
tenants = [t for t in rng.integers(0, spec.n_users, size=8192) if off[t + 1] - off[t] <= 1024][:2048]
fl = [[make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)] for t in tenants]
seg = np.array([int(off[t + 1] - off[t]) for t in tenants])
q1 = torch.from_numpy(synth.query_vector(spec, 1)[None, :]).to(DEV)
q64 = torch.from_numpy(np.stack([synth.query_vector(spec, 100 + i) for i in range(64)])).to(DEV)
def run(q, filters_of, k, n, B):
    oi = torch.empty((B, k), dtype=torch.int64, device=DEV); os_ = torch.empty((B, k), dtype=torch.float64, device=DEV)
    oc = torch.empty(B, dtype=torch.int32, device=DEV)
    for i in range(20): eng.search_device(q, filters_of(i), k, oi, os_, oc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): eng.search_device(q, filters_of(i), k, oi, os_, oc)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for k in (10, 16, 17):
    us = run(q1, lambda i: fl[i % len(fl)], k, 2048, 1)
    print(json.dumps({"batch": 1, "k": k, "us_per_query": round(us, 2), "mean_rows": float(seg.mean()), "median_rows": float(np.median(seg))}), flush=True)
groups = [[fl[(g * 64 + j) % len(fl)][0] for j in range(64)] for g in range(16)]
for k in (10, 17):
    us = run(q64, lambda i: groups[i % 16], k, 300, 64)
    print(json.dumps({"batch": 64, "k": k, "us_per_call": round(us, 2), "qps": round(64e6 / us)}), flush=True)
eng.close()
