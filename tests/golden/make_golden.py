#!/usr/bin/env python
"""Generates the committed golden fixtures (run in the BUILD container, where /root/reference exists;
the GPU box only reads the JSON files this writes).

  retrieve_golden.json   outputs of the REFERENCE's own function body
      /root/reference/tools/qdrant_tool.py::retrieve_transactions, imported UNMODIFIED with stand-in
      modules for its missing third-party imports (langchain_openai, langchain_core, qdrant_client,
      config) and executed over the oracle's model of qdrant local mode (oracle/local_mode.py)
      on a seeded synthetic corpus. What this pins: the tool-level behaviour (filter construction,
      limit rule, post-check, score order, error convention) end to end. What it cannot pin: the
      arithmetic of qdrant-client itself (not installable - parity unpinned).
  groupby_golden.json    outputs of the reference's aggregation expression
      pd.read_json(...).groupby(group_by)[y_axis].sum()  (tools/plot_tool.py:29,45,52), i.e. pandas
      itself, on a small transactions JSON with null keys, negative amounts and a zero-sum group;
      plus the (user, category, month) group-by of a seeded column corpus.
"""
import datetime as real_datetime
import importlib.util
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from financial_chatbot_llm_b200 import qdrant_models, synth  # noqa: E402
from oracle import groupby as ogroupby  # noqa: E402
from oracle import local_mode  # noqa: E402

REF_TOOL = "/root/reference/tools/qdrant_tool.py"
FIXED_NOW = real_datetime.datetime(2023, 12, 20, 12, 0, 0)
SPEC = dict(n_rows=3000, dim=128, n_users=12, seed=2024)


class _Resp:
    def __init__(self, points):
        self.points = points


class _Hit:
    def __init__(self, d):
        self.id, self.score, self.payload, self.version, self.vector = d["id"], d["score"], d["payload"], 0, None


class OracleClient:
    """qdrant_client.QdrantClient stand-in backed by the oracle's local-mode model."""
    collection = None

    def __init__(self, url=None, api_key=None, **kw):
        pass

    def query_points(self, collection_name, query, limit=10, search_params=None, query_filter=None, **kw):
        assert collection_name == "transactions"
        return _Resp([_Hit(d) for d in OracleClient.collection.query_points(query, query_filter, limit=limit)])


def load_reference_tool(embed):
    def tool(*a, **kw):
        if a and callable(a[0]) and not kw:
            return a[0]
        return lambda fn: fn
    mods = {
        "langchain_openai": types.SimpleNamespace(OpenAIEmbeddings=lambda model=None: types.SimpleNamespace(embed_query=embed)),
        "langchain_core": types.ModuleType("langchain_core"),
        "langchain_core.tools": types.SimpleNamespace(tool=tool),
        "qdrant_client": types.SimpleNamespace(QdrantClient=OracleClient),
        "qdrant_client.http": types.SimpleNamespace(models=qdrant_models),
        "qdrant_client.http.models": qdrant_models,
        "config": types.SimpleNamespace(
            get_logger=lambda name: __import__("logging").getLogger(name), QDRANT_URL="http://unused",
            QDRANT_API_KEY="unused", QDRANT_COLLECTION_NAME="transactions", OPENAI_EMBEDDINGS_MODEL_NAME="unused"),
    }
    saved = {k: sys.modules.get(k) for k in mods}
    sys.modules.update(mods)
    try:
        spec = importlib.util.spec_from_file_location("ref_qdrant_tool", REF_TOOL)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v

    class FixedDatetime(real_datetime.datetime):
        @classmethod
        def now(cls, tz=None):
            return FIXED_NOW
    mod.datetime = FixedDatetime            # the tool calls datetime.now() (qdrant_tool.py:117)
    return mod


def retrieve_cases(hc):
    u = hc.users
    return [
        dict(user_id=u[0], num_transactions=10, time_period_days=None, search_query="q0"),
        dict(user_id=u[3], num_transactions=5, time_period_days=90, search_query="q1"),
        dict(user_id=u[7], num_transactions=16, time_period_days=365, search_query="q2"),
        dict(user_id=u[11], num_transactions=None, time_period_days=None, search_query="q3"),     # limit 10000
        dict(user_id=u[2], num_transactions=200, time_period_days=500, search_query="q4"),
        dict(user_id=u[5], num_transactions=3, time_period_days=0, search_query="q5"),            # 0 => no date filter
        dict(user_id="", num_transactions=10, time_period_days=None, search_query="q6"),          # guard
        dict(user_id="user_unknown", num_transactions=10, time_period_days=None, search_query="q7"),
        dict(user_id=u[1], num_transactions=10, time_period_days=1, search_query="q8"),           # nothing recent
    ]


def make_retrieve():
    spec = synth.SynthSpec(**SPEC)
    hc = synth.make_host_corpus(spec)
    col = local_mode.LocalCollection(spec.dim)
    col.upsert([int(i) for i in hc.ids], hc.vectors, [hc.payload(r) for r in range(spec.n_rows)])
    OracleClient.collection = col
    embed = lambda text: synth.query_vector(spec, int(text[1:])).tolist()
    mod = load_reference_tool(embed)
    out = []
    for case in retrieve_cases(hc):
        out.append({"args": case, "result": mod.retrieve_transactions(**case)})
    return {"spec": SPEC, "fixed_now": FIXED_NOW.isoformat(), "reference_file": "tools/qdrant_tool.py",
            "note": "stored vectors are the synthetic rows as generated (the engine and the oracle both "
                    "score cosine, so upsert-time normalisation does not change ids or order)",
            "cases": out}


def make_groupby():
    recs = []
    cats = ["Groceries", "Dining", "Rent", None, "Fuel", "Refunds"]
    rng = np.random.default_rng(7)
    n_ref = 0
    for i in range(240):
        c = cats[int(rng.integers(0, len(cats)))]
        cents = int(rng.integers(-50000, 200000))
        if c == "Refunds":
            cents = 1250 if n_ref % 2 == 0 else -1250
            n_ref += 1
        recs.append({"date": f"2024-{1 + i % 12:02d}-{1 + i % 28:02d}", "category": c, "merchant": f"m{i % 17}",
                     "amount_cents": cents, "amount": cents / 100.0, "qty": int(rng.integers(1, 5))})
    if n_ref % 2:
        recs.append({"date": "2024-01-01", "category": "Refunds", "merchant": "m0", "amount_cents": -1250,
                     "amount": -12.5, "qty": 1})
    tj = json.dumps(recs)
    res = {}
    for gb, y in [("category", "amount_cents"), ("category", "amount"), ("merchant", "qty"), ("qty", "amount_cents")]:
        s = ogroupby.groupby_sum_json(tj, gb, y)
        res[f"{gb}|{y}"] = {"index": [x if isinstance(x, str) else int(x) for x in s.index.tolist()],
                            "values": [float(v) if s.dtype.kind == "f" else int(v) for v in s.tolist()],
                            "dtype": str(s.dtype)}
    spec = synth.SynthSpec(n_rows=50000, dim=8, n_users=23, seed=99)
    off = synth.segment_offsets(spec)
    d, c, v = synth.dates(spec, 0, spec.n_rows), synth.categories(spec, 0, spec.n_rows), synth.cents(spec, 0, spec.n_rows)
    ser = ogroupby.groupby_user_category_month(synth.row_users(off, np.arange(spec.n_rows)), c, d, v)
    idx = ser.index.to_frame(index=False)
    return {"transactions_json": tj, "json_groupby": res,
            "columns": {"spec": dict(n_rows=50000, dim=8, n_users=23, seed=99), "user": idx["user"].tolist(),
                        "category": idx["category"].tolist(), "month": idx["month"].tolist(),
                        "sum": [int(x) for x in ser.tolist()], "pandas": __import__("pandas").__version__}}


if __name__ == "__main__":
    with open(os.path.join(HERE, "retrieve_golden.json"), "w") as f:
        json.dump(make_retrieve(), f, indent=1)
    with open(os.path.join(HERE, "groupby_golden.json"), "w") as f:
        json.dump(make_groupby(), f)
    print("wrote golden fixtures")
