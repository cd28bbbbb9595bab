"""Run ONE secondary workload a few times (for ncu): python scripts/run_one.py cfg3|cfg5|cfg5s|cfg2b|k10000 [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import __graft_entry__ as g
g.build()
from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter, synth_columns
which = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
DEV = torch.device("cuda", 0)
if which == "cfg3":
    rows, dim, B = 10_000_000, 1536, 256
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=3, n_users=1)
    eng = Engine(device=0, max_batch=256, max_k=16)
    loader.collection_synthetic(spec, eng, bf16=True, with_cents=False)
    q = torch.from_numpy(np.stack([synth.query_vector(spec, 500 + i) for i in range(B)])).to(DEV)
    oi = torch.empty((B, 10), dtype=torch.int64, device=DEV); os_ = torch.empty((B, 10), dtype=torch.float64, device=DEV)
    oc = torch.empty(B, dtype=torch.int32, device=DEV)
    from financial_chatbot_llm_b200 import _lib
    prof = torch.zeros(10, dtype=torch.int64, device=DEV)
    _lib.load().fcb_debug_dense_prof(prof.data_ptr())
    for i in range(steps):
        eng.search_device(q, [make_filter(0, rows)] * B, 10, oi, os_, oc)
    torch.cuda.synchronize()
    names = ["producer total", "  wait free stage", "mma total", "  wait tmem drained", "  wait full stage", "epilogue total",
             "  wait accumulators", "  scan tmem", "  fold candidates", "super-tiles"]
    for n, v in zip(names, prof.cpu().tolist()): print(f"{n:24s} {v}")
elif which in ("cfg5", "cfg5s"):
    n, users = 1_000_000_000, 100_000
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=users, seed=5, zipf_a=1.1)
    d = torch.empty(n, dtype=torch.int32, device=DEV); c = torch.empty(n, dtype=torch.uint8, device=DEV)
    v = torch.empty(n, dtype=torch.int32, device=DEV)
    synth_columns(0, n, spec.seed, 16, spec.date_span_seconds, dates=d, categories=c, cents=v)
    off = synth.segment_offsets(spec)
    if which == "cfg5s":
        from financial_chatbot_llm_b200.engine import synth_dates_sorted
        synth_dates_sorted(d, torch.from_numpy(off).to(DEV), 0, spec.date_span_seconds)
    eng = Engine(device=0, max_batch=1, max_k=1)
    eng.load_shard(torch.zeros((n, 4), dtype=torch.float32, device=DEV), dates=d, categories=c, cents=v,
                   seg_offsets=torch.from_numpy(off).to(DEV))
    for i in range(steps):
        eng.aggregate(0, users, (2022 - 1970) * 12, 24, 16)
    torch.cuda.synchronize()
elif which == "k10000":
    rows, dim, k = 10_000_000, 1536, 10000
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=1, n_users=1)
    eng = Engine(device=0, max_batch=1, max_k=k)
    loader.collection_synthetic(spec, eng, with_cents=False)
    q = torch.from_numpy(synth.query_vector(spec, 1)[None, :]).to(DEV)
    oi = torch.empty((1, k), dtype=torch.int64, device=DEV); os_ = torch.empty((1, k), dtype=torch.float64, device=DEV)
    oc = torch.empty(1, dtype=torch.int32, device=DEV)
    for i in range(steps):
        eng.search_device(q, [make_filter(0, rows, synth.DATE_BASE, None)], k, oi, os_, oc)
    torch.cuda.synchronize()
elif which == "cfg2b":
    rows, dim = 10_000_000, 1536
    spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=2, n_users=10000, zipf_a=1.1)
    eng = Engine(device=0, max_batch=64, max_k=16)
    col = loader.collection_synthetic(spec, eng, with_cents=False)
    off = col.offsets
    q = torch.from_numpy(synth.query_vector(spec, 1)[None, :]).to(DEV)
    oi = torch.empty((1, 10), dtype=torch.int64, device=DEV); os_ = torch.empty((1, 10), dtype=torch.float64, device=DEV)
    oc = torch.empty(1, dtype=torch.int32, device=DEV)
    for i in range(steps):
        t = 5000 + i
        eng.search_device(q, [make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)], 10, oi, os_, oc)
    torch.cuda.synchronize()
print("done", which)
