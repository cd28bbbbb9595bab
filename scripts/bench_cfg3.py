#!/usr/bin/env python
"""cfg3 only (10M x 1536 bf16, shared segment, batch 256/128/64): one line per batch size. For A/B runs of two
builds on the same box: FCB_LIB_PATH=<other libfcb200.so> python scripts/bench_cfg3.py"""
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
rows, dim, K = 10_000_000, 1536, 10
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
batches = [int(b) for b in sys.argv[2].split(",")] if len(sys.argv) > 2 else [256, 128, 64]
spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=3, n_users=1)
eng = Engine(device=0, max_batch=256, max_k=16)
loader.collection_synthetic(spec, eng, bf16=True, with_cents=False)
oi = torch.empty((256, K), dtype=torch.int64, device=DEV); os_ = torch.empty((256, K), dtype=torch.float64, device=DEV)
oc = torch.empty(256, dtype=torch.int32, device=DEV)
for B in batches:
    q = torch.from_numpy(np.stack([synth.query_vector(spec, 500 + i) for i in range(B)])).to(DEV)
    f = [make_filter(0, rows)] * B
    for _ in range(5):
        eng.search_device(q, f, K, oi, os_, oc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.search_device(q, f, K, oi, os_, oc)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(json.dumps({"lib": os.environ.get("FCB_LIB_PATH", "in-tree"), "batch": B, "ms": round(ms, 4),
                      "tensor_TFLOPs": round(2.0 * rows * dim * (((B + 31) // 32) * 32) / ms / 1e9, 1),
                      "redo": eng.last_dense_redo(B), "widened": eng.last_dense_widened(B), "exact_reruns": eng.exact_reruns, "ids_checksum": int(oi[:B].sum().item())}), flush=True)
eng.close()
