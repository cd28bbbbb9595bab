"""The reference's own tool file, UNMODIFIED, imported over the compat shim (CPU; needs /root/reference,
which exists in the build container only — skipped elsewhere). qdrant_client resolves to our stand-in via
compat.install_qdrant_shim(); the other missing third-party imports of that file (langchain, its config
module) are stubbed here in the test. The backend behind the qdrant-shaped client is the oracle's
local-mode model, so this checks the BOUNDARY (constructor keywords, models, query_points keywords,
.points[i].payload) with the reference's code as the caller."""
import datetime as dt
import importlib.util
import json
import os
import sys
import types

import numpy as np
import pytest

REF_TOOL = "/root/reference/tools/qdrant_tool.py"
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(not os.path.exists(REF_TOOL), reason="reference tree not present on this machine")
def test_reference_tool_file_runs_unmodified_over_the_shim(monkeypatch):
    from financial_chatbot_llm_b200 import compat, synth
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.qdrant_models import QueryResponse, ScoredPoint
    from oracle import local_mode

    g = json.load(open(os.path.join(HERE, "golden", "retrieve_golden.json")))
    spec = synth.SynthSpec(**g["spec"])
    hc = synth.make_host_corpus(spec)
    col = local_mode.LocalCollection(spec.dim)
    col.upsert([int(i) for i in hc.ids], hc.vectors, [hc.payload(r) for r in range(spec.n_rows)])

    class OracleBacked(B200QdrantClient):
        """Same class the shim hands out, with the CPU oracle standing in for the GPU engine (test only)."""
        def query_points(self, collection_name, query=None, query_filter=None, search_params=None, limit=10, **kw):
            assert collection_name == "transactions" and search_params.hnsw_ef == 128 and search_params.exact is False
            return QueryResponse(points=[ScoredPoint(**d) for d in col.query_points(query, query_filter, limit=limit)])

    def tool(*a, **kw):
        return a[0] if a and callable(a[0]) and not kw else (lambda fn: fn)
    for name, mod in {
        "langchain_openai": types.SimpleNamespace(OpenAIEmbeddings=lambda model=None: types.SimpleNamespace(
            embed_query=lambda text: synth.query_vector(spec, int(text[1:])).tolist())),
        "langchain_core": types.ModuleType("langchain_core"),
        "langchain_core.tools": types.SimpleNamespace(tool=tool),
        "config": types.SimpleNamespace(get_logger=lambda n: __import__("logging").getLogger(n), QDRANT_URL="http://unused",
                                        QDRANT_API_KEY="k", QDRANT_COLLECTION_NAME="transactions",
                                        OPENAI_EMBEDDINGS_MODEL_NAME="unused"),
    }.items():
        monkeypatch.setitem(sys.modules, name, mod)
    for name in ("qdrant_client", "qdrant_client.http", "qdrant_client.http.models"):
        monkeypatch.setitem(sys.modules, name, sys.modules.get(name))     # restored / removed after the test
    compat.install_qdrant_shim(client_factory=OracleBacked, force=True)

    spec_ = importlib.util.spec_from_file_location("ref_qdrant_tool_dropin", REF_TOOL)
    mod = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(mod)                                         # the reference file, as it is
    assert isinstance(mod.qdrant_client.get_client(), B200QdrantClient)   # QdrantHttpClient(url=, api_key=) accepted

    now = dt.datetime.fromisoformat(g["fixed_now"])

    class FixedDatetime(dt.datetime):
        @classmethod
        def now(cls, tz=None):
            return now
    mod.datetime = FixedDatetime
    for case in g["cases"]:
        assert mod.retrieve_transactions(**case["args"]) == case["result"], case["args"]
    for name in ("qdrant_client", "qdrant_client.http", "qdrant_client.http.models"):
        if sys.modules.get(name) is None:
            sys.modules.pop(name, None)
