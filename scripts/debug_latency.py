"""Phase timeline of the fused scan on small tenant segments (diagnostic)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from financial_chatbot_llm_b200 import _lib, loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter
rows, dim = 2_000_000, 1536
spec = synth.SynthSpec(n_rows=rows, dim=dim, seed=2, n_users=2000, zipf_a=1.1)
eng = Engine(device=0, max_batch=1, max_k=16)
col = loader.collection_synthetic(spec, eng, with_cents=False)
off = col.offsets
q = torch.from_numpy(synth.query_vector(spec, 1)[None, :]).cuda()
oi = torch.empty((1, 10), dtype=torch.int64, device="cuda"); os_ = torch.empty((1, 10), dtype=torch.float64, device="cuda")
oc = torch.empty(1, dtype=torch.int32, device="cuda")
tb = torch.zeros(9, dtype=torch.int64, device="cuda")
lib = _lib.load()
names = ["start", "query staged", "scan done", "cta list written", "last cta starts", "lists merged", "rescore starts", "rescore ends", "written"]
for t in (1500, 900, 400, 100, 20, 3):
    n = int(off[t + 1] - off[t])
    f = [make_filter(int(off[t]), int(off[t + 1]), synth.DATE_BASE, None)]
    acc = np.zeros(9)
    lib.fcb_debug_scan_timing(tb.data_ptr())
    for i in range(20):
        eng.search_device(q, f, 10, oi, os_, oc)
        torch.cuda.synchronize()
        v = tb.cpu().numpy().astype(np.float64)
        if i >= 5: acc += v - v[0]
    lib.fcb_debug_scan_timing(None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(200): eng.search_device(q, f, 10, oi, os_, oc)
    e1.record(); torch.cuda.synchronize()
    print(f"tenant {t}: {n} rows; back-to-back {e0.elapsed_time(e1) / 200 * 1e3:.1f} us/query; phases (us since start): " +
          ", ".join(f"{nm} {a / 15 / 1e3:.2f}" for nm, a in zip(names, acc)))
