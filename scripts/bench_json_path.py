#!/usr/bin/env python
"""p50 of the tool-shaped aggregation call (tools/plot_tool.py:29,45,52: read_json -> groupby(key)[value].sum()) through
aggregate.grouped_sum_json (the reduction on the GPU, fcb_aggregate_columns) beside the same call done by pandas alone, for the
sizes an LLM tool call carries (10^2..10^4 transactions). JSON parsing is part of both."""
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import pandas as pd

import __graft_entry__ as graft
from financial_chatbot_llm_b200 import aggregate

graft.build()
rng = np.random.default_rng(4)
cats = ["groceries", "rent", "salary", "travel", "dining", "utilities", "health", "fun"]
for n in (100, 1000, 10000):
    rows = [{"date": "2023-%02d-%02d" % (rng.integers(1, 13), rng.integers(1, 28)), "category": cats[int(rng.integers(0, len(cats)))],
             "amount": round(float(rng.integers(-500000, 500000)) / 100.0, 2)} for _ in range(n)]
    txt = json.dumps(rows)
    want = pd.read_json(io.StringIO(txt)).groupby("category")["amount"].sum()
    got = aggregate.grouped_sum_json(txt, "category", "amount")
    assert got.index.tolist() == want.index.tolist() and np.allclose(got.to_numpy(), want.to_numpy(), rtol=0, atol=1e-6)

    def p50(fn, reps=60):
        for _ in range(5):
            fn()
        t = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); t.append(time.perf_counter() - t0)
        return 1e3 * float(np.median(t))
    print(json.dumps({"workload": f"tool-shaped group-by of {n} transactions (JSON text in, Series out), group_by=category, y_axis=amount",
                      "engine_p50_ms": p50(lambda: aggregate.grouped_sum_json(txt, "category", "amount")),
                      "pandas_p50_ms": p50(lambda: pd.read_json(io.StringIO(txt)).groupby("category")["amount"].sum()),
                      "read_json_alone_p50_ms": p50(lambda: pd.read_json(io.StringIO(txt))),
                      "equal_to_pandas": True}), flush=True)
