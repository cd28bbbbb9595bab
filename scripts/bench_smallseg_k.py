#!/usr/bin/env python
"""The tool's DEFAULT call shape: limit = 10000 (tools/qdrant_tool.py:145) against ordinary tenants (10k Zipf tenants over
10M x 1536 fp32, median segment 128 rows): every row of the tenant comes back, ranked. One line per k."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter

DEV = torch.device("cuda", 0)
rows, dim = 10_000_000, 1536
spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=2, n_users=10000, zipf_a=1.1)
eng = Engine(device=0, max_batch=64, max_k=10000)
col = loader.collection_synthetic(spec, eng, with_cents=False)
off = col.offsets
rng = np.random.default_rng(0)
tenants = rng.integers(0, spec.n_users, size=2048)
fl = [[make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)] for t in tenants]
seg = np.array([int(off[t + 1] - off[t]) for t in tenants])
q = torch.from_numpy(synth.query_vector(spec, 1)[None, :]).to(DEV)
for k in (10, 100, 10000):
    oi = torch.empty((1, k), dtype=torch.int64, device=DEV); os_ = torch.empty((1, k), dtype=torch.float64, device=DEV)
    oc = torch.empty(1, dtype=torch.int32, device=DEV)
    for i in range(50):
        eng.search_device(q, fl[i], k, oi, os_, oc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(2048):
        eng.search_device(q, fl[i], k, oi, os_, oc)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 2048 * 1e3
    print(json.dumps({"workload": "10k Zipf tenants, batch 1, uniform random tenant", "k": k, "us_per_query": round(us, 2),
                      "qps": round(1e6 / us), "launches_per_query": eng.last_search_launches,
                      "mean_segment_rows": float(seg.mean()), "median_segment_rows": float(np.median(seg)),
                      "max_segment_rows": int(seg.max())}), flush=True)
# host API wall clock (H2D query, launch, D2H results, sync), batch 1
import time
qh = synth.query_vector(spec, 5)[None, :]
for k in (10, 10000):
    for i in range(30):
        eng.search_host(qh, fl[i], k)
    lat = []
    for i in range(1000):
        t1 = time.perf_counter(); eng.search_host(qh, fl[i], k); lat.append(time.perf_counter() - t1)
    print(json.dumps({"workload": "host API (Engine.search_host) wall clock, batch 1", "k": k, "p50_us": round(1e6 * float(np.median(lat)), 1),
                      "mean_us": round(1e6 * float(np.mean(lat)), 1), "p99_us": round(1e6 * float(np.percentile(lat, 99)), 1)}), flush=True)
# 64 default-limit calls per launch (what the serving layer's micro-batching produces), tenants of at most 1024 rows
small = [f[0] for f, n in zip(fl, seg) if n <= 1024]
q64 = torch.from_numpy(np.stack([synth.query_vector(spec, 100 + i) for i in range(64)])).to(DEV)
k = 10000
oi = torch.empty((64, k), dtype=torch.int64, device=DEV); os_ = torch.empty((64, k), dtype=torch.float64, device=DEV)
oc = torch.empty(64, dtype=torch.int32, device=DEV)
groups = [small[(g * 64) % (len(small) - 64):][:64] for g in range(16)]
for g in groups[:4]:
    eng.search_device(q64, g, k, oi, os_, oc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(200):
    eng.search_device(q64, groups[i % 16], k, oi, os_, oc)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 200 * 1e3
print(json.dumps({"workload": "64 default-limit calls (k = 10000) per launch, tenants <= 1024 rows", "us_per_call": round(us, 2),
                  "qps": round(64e6 / us), "launches_per_call": eng.last_search_launches}), flush=True)
# one big-but-ordinary tenant: the largest below 10000 rows
sizes = np.diff(off)
cand = np.flatnonzero(sizes < 10000)
t = int(cand[np.argmax(sizes[cand])])
f1 = [make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)]
k = 10000
oi = torch.empty((1, k), dtype=torch.int64, device=DEV); os_ = torch.empty((1, k), dtype=torch.float64, device=DEV)
oc = torch.empty(1, dtype=torch.int32, device=DEV)
for i in range(20):
    eng.search_device(q, f1, k, oi, os_, oc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(200):
    eng.search_device(q, f1, k, oi, os_, oc)
e1.record()
torch.cuda.synchronize()
print(json.dumps({"workload": f"one tenant of {int(sizes[t])} rows, k = 10000 (everything returned, ranked)",
                  "us_per_query": round(e0.elapsed_time(e1) / 200 * 1e3, 2), "count": int(oc.item())}), flush=True)
eng.close()
