"""Golden fixtures (tests/golden/*.json, made by tests/golden/make_golden.py from the reference's own
function body and from pandas) against the oracle and the host mirror. CPU only."""
import datetime as dt
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def load(name):
    with open(os.path.join(HERE, "golden", name)) as f:
        return json.load(f)


class OracleBackedClient:
    """qdrant-shaped client over the oracle's local-mode model (test only)."""
    def __init__(self, col):
        self.col = col

    def query_points(self, collection_name, query, limit=10, search_params=None, query_filter=None, **kw):
        from financial_chatbot_llm_b200.qdrant_models import QueryResponse, ScoredPoint
        return QueryResponse(points=[ScoredPoint(**d) for d in self.col.query_points(query, query_filter, limit=limit)])


@pytest.fixture(scope="module")
def retrieve_env():
    from financial_chatbot_llm_b200 import synth
    from oracle import local_mode
    g = load("retrieve_golden.json")
    spec = synth.SynthSpec(**g["spec"])
    hc = synth.make_host_corpus(spec)
    col = local_mode.LocalCollection(spec.dim)
    col.upsert([int(i) for i in hc.ids], hc.vectors, [hc.payload(r) for r in range(spec.n_rows)])
    return g, spec, hc, col


def test_host_mirror_reproduces_the_reference_function(retrieve_env):
    """tools.retrieve_transactions (our mirror) == /root/reference/tools/qdrant_tool.py's function
    (golden), both over the same oracle backend: the mirror's host logic is the reference's."""
    from financial_chatbot_llm_b200 import synth, tools
    g, spec, hc, col = retrieve_env
    now = dt.datetime.fromisoformat(g["fixed_now"])
    tools.configure(OracleBackedClient(col), lambda text: synth.query_vector(spec, int(text[1:])).tolist(),
                    now=lambda: now)
    for case in g["cases"]:
        assert tools.retrieve_transactions(**case["args"]) == case["result"], case["args"]
    assert [len(c["result"]) for c in g["cases"]][6:8] == [0, 0]


def test_c_restatement_reproduces_the_golden_order(retrieve_env):
    """The pinned-order C oracle yields the same page_content sequence as the golden (which came
    from the numpy model): the two restatements agree at tool level too."""
    from financial_chatbot_llm_b200 import synth
    from oracle import canon
    g, spec, hc, col = retrieve_env
    now = dt.datetime.fromisoformat(g["fixed_now"])
    for case in g["cases"]:
        a = case["args"]
        if a["user_id"] not in hc.users:
            continue
        u = hc.users.index(a["user_id"])
        lo = int((now - dt.timedelta(days=a["time_period_days"])).timestamp()) if a["time_period_days"] else None
        k = a["num_transactions"] if a["num_transactions"] is not None else 10000
        ids, _ = canon.search(hc.vectors, synth.query_vector(spec, int(a["search_query"][1:])), k,
                              row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]), dates=hc.dates, date_lo=lo)
        assert [hc.payload(int(r))["page_content"] for r in ids] == case["result"]


def test_pandas_groupby_golden_is_reproducible():
    from oracle import groupby
    g = load("groupby_golden.json")
    for key, want in g["json_groupby"].items():
        gb, y = key.split("|")
        s = groupby.groupby_sum_json(g["transactions_json"], gb, y)
        assert [x if isinstance(x, str) else int(x) for x in s.index.tolist()] == want["index"]
        assert np.allclose(s.to_numpy().astype(float), want["values"], rtol=0, atol=1e-9) and str(s.dtype) == want["dtype"]
    assert g["json_groupby"]["category|amount_cents"]["values"][g["json_groupby"]["category|amount_cents"]["index"].index("Refunds")] == 0


def test_c_groupby_reproduces_pandas_golden_columns():
    from financial_chatbot_llm_b200 import synth
    from oracle import canon
    c = load("groupby_golden.json")["columns"]
    spec = synth.SynthSpec(**c["spec"])
    off = synth.segment_offsets(spec)
    u, cat, mo, s, _ = canon.groupby(off, synth.dates(spec, 0, spec.n_rows), synth.categories(spec, 0, spec.n_rows),
                                     synth.cents(spec, 0, spec.n_rows))
    assert u.tolist() == c["user"] and cat.tolist() == c["category"] and mo.tolist() == c["month"] and s.tolist() == c["sum"]
