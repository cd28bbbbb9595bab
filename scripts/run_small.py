"""A few single queries against small tenants (for an ncu launch list): python scripts/run_small.py <k> [n]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from financial_chatbot_llm_b200 import loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter
DEV = torch.device("cuda", 0)
k = int(sys.argv[1]); n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
spec = synth.SynthSpec(n_rows=2_000_000, dim=1536, seed=2, n_users=2000, zipf_a=1.1)
eng = Engine(device=0, max_batch=1, max_k=10000)
col = loader.collection_synthetic(spec, eng, with_cents=False)
off = col.offsets
ten = [t for t in range(spec.n_users) if 64 <= off[t + 1] - off[t] <= 1024][-n:]
q = torch.from_numpy(synth.query_vector(spec, 1)[None, :]).to(DEV)
oi = torch.empty((1, k), dtype=torch.int64, device=DEV); os_ = torch.empty((1, k), dtype=torch.float64, device=DEV)
oc = torch.empty(1, dtype=torch.int32, device=DEV)
for t in ten:
    eng.search_device(q, [make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)], k, oi, os_, oc)
torch.cuda.synchronize()
print("rows", [int(off[t + 1] - off[t]) for t in ten][:10])
