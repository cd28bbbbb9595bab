#!/usr/bin/env python
"""Served path: ONE EngineServer process owns the GPU-resident collection, N worker processes (the
reference runs 3-6 gunicorn workers, gunicorn.conf.py:8 / Dockerfile:39, one query in flight each) call it through
EngineClient.query_points over the Unix-domain socket. Reports queries/s, client-side p50 / p99 latency and the server's
micro-batch histogram. Corpus: 10k Zipf tenants x 1536 fp32 (BASELINE configs[1] shape; --rows to shrink).
  python scripts/bench_serving.py [--workers 6] [--queries 2000] [--rows 10000000] [--window-us 200]"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def worker(addr, wid, n_queries, n_users, dim, seed, out):
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200 import serving, synth
    spec = synth.SynthSpec(n_rows=n_users, dim=dim, seed=seed, n_users=n_users)
    users = synth.user_ids(n_users)
    rng = np.random.default_rng(1000 + wid)
    qs = [synth.query_vector(spec, 10_000 * wid + i).tolist() for i in range(64)]
    cl = serving.EngineClient(address=addr)
    flt = lambda u: m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=users[u]))])
    for i in range(20):
        cl.query_points("transactions", query=qs[i % 64], limit=10, query_filter=flt(int(rng.integers(n_users))))
    lat = []
    t0 = time.perf_counter()
    for i in range(n_queries):
        u = int(rng.integers(n_users))
        t1 = time.perf_counter()
        pts = cl.query_points("transactions", query=qs[i % 64], limit=10, query_filter=flt(u)).points
        lat.append(time.perf_counter() - t1)
        assert all(p.payload["metadata"]["user_id"] == users[u] for p in pts)
    out.put((wid, time.perf_counter() - t0, lat))
    cl.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=6)
    ap.add_argument("--queries", type=int, default=2000)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--users", type=int, default=10_000)
    ap.add_argument("--window-us", type=float, default=200.0)
    a = ap.parse_args()
    import torch
    import __graft_entry__ as graft
    graft.build()
    from financial_chatbot_llm_b200 import loader, serving, synth
    from financial_chatbot_llm_b200.engine import Engine
    dim, seed = 1536, 2
    spec = synth.SynthSpec(n_rows=a.rows, dim=dim, seed=seed, n_users=a.users, zipf_a=1.1)
    eng = Engine(device=0, max_batch=64, max_k=16)
    col = loader.collection_synthetic(spec, eng, with_cents=False)
    addr = os.path.join(tempfile.mkdtemp(), "engine.sock")
    for workers in sorted({1, a.workers}):
        srv = serving.EngineServer(col, addr, window_us=a.window_us).start()
        ctx = mp.get_context("spawn")
        out = ctx.Queue()
        procs = [ctx.Process(target=worker, args=(addr, w, a.queries, a.users, dim, seed, out)) for w in range(workers)]
        t0 = time.perf_counter()
        [p.start() for p in procs]
        res = [out.get(timeout=600) for _ in procs]
        [p.join() for p in procs]
        wall = max(r[1] for r in res)
        lat = np.sort(np.concatenate([np.array(r[2]) for r in res]))
        hist = dict(sorted(srv.stats["batch_hist"].items()))
        print(json.dumps({"workload": f"served path: {workers} worker processes -> 1 EngineServer, {a.rows} x {dim} fp32, {a.users} Zipf tenants, "
                          f"top-10 + payloads, window {a.window_us} us",
                          "workers": workers, "queries": workers * a.queries, "queries_per_s": workers * a.queries / wall,
                          "p50_us": 1e6 * float(lat[len(lat) // 2]), "p99_us": 1e6 * float(lat[int(len(lat) * 0.99)]),
                          "launches": srv.stats["launches"], "max_micro_batch": srv.stats["max_batch"], "micro_batch_hist": hist}), flush=True)
        srv.stop()
    eng.close()


if __name__ == "__main__":
    main()
