#!/usr/bin/env python
"""Latency of ONE small-tenant query through the host entry point (fcb_search_host: host query in, host results out, wall
clock) beside the device time of the same call (fcb_last_search_ms: CUDA events around the kernels). 2M x 1536 fp32, 10k Zipf
tenants, user_id + date filter, top-10."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

import __graft_entry__ as graft
from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter

graft.build()
rows, dim, K = 2_000_000, 1536, 10
spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=2, n_users=10000, zipf_a=1.1)
eng = Engine(device=0, max_batch=4, max_k=16)
col = loader.collection_synthetic(spec, eng, with_cents=False)
off = col.offsets
rng = np.random.default_rng(0)
tenants = rng.integers(0, spec.n_users, size=2048)
fl = [[make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)] for t in tenants]
qh = synth.query_vector(spec, 5)[None, :]
for i in range(50):
    eng.search_host(qh, fl[i], K)
lat, dev = [], []
t0 = time.perf_counter()
for i in range(2000):
    t1 = time.perf_counter()
    eng.search_host(qh, fl[i % 2048], K)
    lat.append(time.perf_counter() - t1)
    if i % 8 == 0:
        dev.append(eng.last_search_ms)
wall = time.perf_counter() - t0
print(json.dumps({"workload": "one small-tenant query through fcb_search_host (wall clock, host in / host out)",
                  "FCB_HOST_DIRECT": os.environ.get("FCB_HOST_DIRECT", "default (1)"), "qps": 2000 / wall,
                  "p50_us": 1e6 * float(np.median(lat)), "p99_us": 1e6 * float(np.percentile(lat, 99)),
                  "device_p50_us": 1e3 * float(np.median(dev)), "median_segment_rows": float(np.median(np.diff(off)))}), flush=True)
eng.close()
