"""Diagnostic: dump the tcgen05 accumulators of the first rows and compare with numpy."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import __graft_entry__ as g
g.build()
from financial_chatbot_llm_b200 import _lib, loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter

dim = int(sys.argv[1]) if len(sys.argv) > 1 else 128
spec = synth.SynthSpec(n_rows=70000, dim=dim, n_users=2, seed=51, whale_rows=69000)
hc = synth.make_host_corpus(spec, bf16=True)
eng = Engine(device=0, max_batch=256, max_k=16)
loader.collection_from_host(hc, eng, bf16=True)
B = 32
qs = np.stack([synth.query_vector(spec, 1000 + i) for i in range(B)])
R = 1024
dbg = torch.zeros((R, 256), dtype=torch.float32, device="cuda")
lib = _lib.load()
lib.fcb_debug_dense_dump(dbg.data_ptr(), R)
ids, sc, cnt = eng.search_host(qs, [make_filter(0, 69000)] * B, 10)
torch.cuda.synchronize()
print("redo", eng.last_dense_redo(B))
got = dbg.cpu().numpy()
qb = synth.bf16_bits_to_f32(synth.f32_to_bf16_bits(qs.astype(np.float32))).reshape(B, dim).astype(np.float64)
want = hc.vectors[:R].astype(np.float64) @ qb.T
print("want[0,:4]", want[0, :4], "got[0,:4]", got[0, :4])
print("want[1,:4]", want[1, :4], "got[1,:4]", got[1, :4])
print("want[130,:4]", want[130, :4], "got[130,:4]", got[130, :4])
err = np.abs(got[:, :B] - want)
print("max abs err", err.max(), "rows with err>1e-2:", int((err.max(axis=1) > 1e-2).sum()), "of", R)
print("pad cols max", np.abs(got[:, B:]).max())
bad = np.where(err.max(axis=1) > 1e-2)[0]
print("first bad rows", bad[:20])
# does got match some permutation? check row r vs want rows
if len(bad):
    r = bad[0]
    d = np.abs(want[:, :4] - got[r, :4]).sum(axis=1)
    print("row", r, "best matching want row", int(np.argmin(d)), d.min())
