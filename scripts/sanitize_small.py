"""Small end-to-end exercise of every kernel for compute-sanitizer (memcheck): tiny shapes, all paths."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from financial_chatbot_llm_b200 import _lib, aggregate, loader, synth
from financial_chatbot_llm_b200.engine import Engine, make_filter, search_noalloc
from oracle import canon

spec = synth.SynthSpec(n_rows=3001, dim=128, n_users=7, seed=3)
hc = synth.make_host_corpus(spec)
for bf16 in (False, True):
    eng = Engine(device=0, max_batch=8, max_k=200)
    loader.collection_from_host(hc if not bf16 else synth.make_host_corpus(spec, bf16=True), eng, bf16=bf16)
    q = np.stack([synth.query_vector(spec, i) for i in range(5)])
    t0 = int(np.median(hc.dates))
    fl = [make_filter(int(hc.offsets[u]), int(hc.offsets[u + 1]), t0 if u % 2 else None, None, None if u % 3 else 0x0ff0) for u in range(5)]
    for k in (1, 10, 16, 17, 200):
        ids, sc, cnt = eng.search_host(q, fl, k)
    if not bf16:
        oi, os_ = canon.search(hc.vectors, q[0], 10, row_begin=int(hc.offsets[0]), row_end=int(hc.offsets[1]), categories=hc.categories, category_mask=0x0ff0)
        ids, sc, cnt = eng.search_host(q[:1], fl[:1], 10)
        assert ids[0, :cnt[0]].tolist() == oi.tolist()
    mlo = (2022 - 1970) * 12
    s, p, d = eng.aggregate(0, 7, mlo, 24, 16)
    torch.cuda.synchronize()
    assert int(d.item()) == 0
    eng.close()
# ordered-log aggregation + no-alloc entry
off = synth.segment_offsets(spec)
dev = torch.device("cuda", 0)
dsorted = synth.dates_sorted(spec, off, 0, spec.n_rows)
eng = Engine(device=0, max_batch=1, max_k=16)
vec = torch.from_numpy(hc.vectors).to(dev)
dd, cc, vv = torch.from_numpy(dsorted).to(dev), torch.from_numpy(hc.categories).to(dev), torch.from_numpy(hc.cents).to(dev)
eng.load_shard(vec, dates=dd, categories=cc, cents=vv, seg_offsets=torch.from_numpy(off).to(dev))
assert eng.dates_sorted
s, p, d = eng.aggregate(0, 7, (2022 - 1970) * 12, 24, 16)
torch.cuda.synchronize()
shard = _lib.FcbShard(vec.data_ptr(), None, None, dd.data_ptr(), cc.data_ptr(), None, None, spec.n_rows, 0, 0, spec.dim, 0)
oi = torch.empty(16, dtype=torch.int64, device=dev); osc = torch.empty(16, dtype=torch.float64, device=dev); oc = torch.empty(1, dtype=torch.int32, device=dev)
search_noalloc(0, shard, torch.from_numpy(synth.query_vector(spec, 1)).to(dev), make_filter(0, 3001), 10, oi, osc, oc)
torch.cuda.synchronize()
eng.close()
# merge kernel
from financial_chatbot_llm_b200 import sharded
m = sharded.ShardMerger(dev, 1, 3, 10)
m.merge(); torch.cuda.synchronize()
print("sanitize_small ok")
