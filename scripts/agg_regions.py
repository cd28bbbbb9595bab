#!/usr/bin/env python
"""Ordered-log aggregation: per-row cost by user size class (cfg5 corpus: 1B rows, 100k Zipf users, ascending size rank)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from financial_chatbot_llm_b200 import synth
from financial_chatbot_llm_b200.engine import Engine, synth_columns, synth_dates_sorted
DEV = torch.device("cuda", 0)
n, users = 1_000_000_000, 100_000
spec = synth.SynthSpec(n_rows=n, dim=4, n_users=users, seed=5, zipf_a=1.1)
d = torch.empty(n, dtype=torch.int32, device=DEV); c = torch.empty(n, dtype=torch.uint8, device=DEV)
v = torch.empty(n, dtype=torch.int32, device=DEV)
synth_columns(0, n, spec.seed, 16, spec.date_span_seconds, dates=d, categories=c, cents=v)
off = synth.segment_offsets(spec)
synth_dates_sorted(d, torch.from_numpy(off).to(DEV), 0, spec.date_span_seconds)
eng = Engine(device=0, max_batch=1, max_k=1)
eng.load_shard(torch.zeros((n, 4), dtype=torch.float32, device=DEV), dates=d, categories=c, cents=v,
               seg_offsets=torch.from_numpy(off).to(DEV))
sz = np.diff(off)
order_ok = bool(np.all(sz[:-1] >= sz[1:]))
mlo = (2022 - 1970) * 12
def t(a, b, reps=10):
    sums = torch.empty((b - a, 16, 24), dtype=torch.int64, device=DEV); pres = torch.empty((b - a, 16, 24), dtype=torch.uint8, device=DEV)
    drop = torch.empty(1, dtype=torch.int64, device=DEV)
    for _ in range(3): eng.aggregate(a, b, mlo, 24, 16, sums, pres, drop)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): eng.aggregate(a, b, mlo, 24, 16, sums, pres, drop)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
print(json.dumps({"sizes_descending": order_ok, "largest": int(sz.max()), "smallest": int(sz.min())}))
cuts = [0, 1, 10, 100, 1000, 10000, 45000, 100000]
for a, b in zip(cuts[:-1], cuts[1:]):
    rows = int(off[b] - off[a]); ms = t(a, b)
    print(json.dumps({"users": [a, b], "rows": rows, "rows_per_user": [int(sz[b - 1]), int(sz[a])], "ms": round(ms, 4),
                      "ns_per_krow": round(ms * 1e6 / (rows / 1e3), 2), "GBs": round(rows * 9 / ms / 1e6, 1)}), flush=True)
print(json.dumps({"all": round(t(0, users), 4)}))
