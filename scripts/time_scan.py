"""Time the fused scan on a whale corpus (no asserts): python scripts/time_scan.py label [f32|bf16] [dim] [rows] [max_batch] [n_users] [date_filter 0|1]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter
label = sys.argv[1] if len(sys.argv) > 1 else ""
bf16 = len(sys.argv) > 2 and sys.argv[2] == "bf16"
dim = int(sys.argv[3]) if len(sys.argv) > 3 else 1536
rows = int(sys.argv[4]) if len(sys.argv) > 4 else 10_000_000
max_batch = int(sys.argv[5]) if len(sys.argv) > 5 else 1
n_users = int(sys.argv[6]) if len(sys.argv) > 6 else 1
date_filter = (sys.argv[7] != '0') if len(sys.argv) > 7 else True
spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=1, n_users=n_users, zipf_a=0.0 if n_users > 1 else 1.1)
eng = Engine(device=0, max_batch=max_batch, max_k=16)
loader.collection_synthetic(spec, eng, bf16=bf16, with_cents=False)
q = torch.from_numpy(synth.query_vector(spec, 0)[None, :]).cuda()
oi = torch.empty((1, 10), dtype=torch.int64, device="cuda"); os_ = torch.empty((1, 10), dtype=torch.float64, device="cuda")
oc = torch.empty(1, dtype=torch.int32, device="cuda")
f = [make_filter(0, rows, synth.DATE_BASE if date_filter else None, None)]
for i in range(5): eng.search_device(q, f, 10, oi, os_, oc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(30): eng.search_device(q, f, 10, oi, os_, oc)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 30
gb = rows * dim * (2 if bf16 else 4) / 1e9
print(label, "bf16" if bf16 else "f32", dim, "ms/query %.3f" % ms, "GB/s %.0f" % (gb / ms * 1e3), flush=True)
