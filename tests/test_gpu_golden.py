"""The CUDA engine behind the tool-shaped API reproduces the golden outputs of the REFERENCE's own
retrieve_transactions function (tests/golden/retrieve_golden.json)."""
import datetime as dt
import json
import os

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_engine_reproduces_reference_function_golden(built):
    from financial_chatbot_llm_b200 import loader, synth, tools
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.engine import Engine
    g = json.load(open(os.path.join(HERE, "golden", "retrieve_golden.json")))
    spec = synth.SynthSpec(**g["spec"])
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=1, max_k=10000)
    cl = B200QdrantClient(url="unused", api_key="unused")
    cl.attach_collection(loader.collection_from_host(hc, eng))
    now = dt.datetime.fromisoformat(g["fixed_now"])
    tools.configure(cl, lambda text: synth.query_vector(spec, int(text[1:])).tolist(), now=lambda: now)
    for case in g["cases"]:
        assert tools.retrieve_transactions(**case["args"]) == case["result"], case["args"]
    eng.close()


def test_default_limit_on_the_documented_engine(built):
    """INTEGRATION.md's recipe, as written (Engine(device=0, max_batch=1, max_k=10000) / the default), must serve the
    reference's DEFAULT call: num_transactions=None -> limit=10000 (tools/qdrant_tool.py:145). An engine with a smaller
    max_k refuses it, and the tool's blanket except would turn that into [] — so also pin that the refusal is loud at
    the client level."""
    from financial_chatbot_llm_b200 import _lib, loader, synth, tools
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.engine import Engine
    from oracle import canon
    spec = synth.SynthSpec(n_rows=9000, dim=128, n_users=5, seed=31)
    hc = synth.make_host_corpus(spec)
    for eng in (Engine(device=0, max_batch=1, max_k=10000), Engine()):
        cl = B200QdrantClient(url="unused", api_key="unused")
        cl.attach_collection(loader.collection_from_host(hc, eng))
        tools.configure(cl, lambda text: synth.query_vector(spec, int(text[1:])).tolist())
        for u in (0, 3):
            got = tools.retrieve_transactions(user_id=hc.users[u], search_query="q7")          # num_transactions=None
            lo, hi = int(hc.offsets[u]), int(hc.offsets[u + 1])
            oi, _ = canon.search(hc.vectors, synth.query_vector(spec, 7), 10000, row_begin=lo, row_end=hi, ids=hc.ids)
            assert len(got) == min(10000, hi - lo) > 0
            assert got == [hc.payload(int(i))["page_content"] for i in oi]
        eng.close()
    small = Engine(device=0, max_batch=1, max_k=16)
    cl = B200QdrantClient(url="unused")
    cl.attach_collection(loader.collection_from_host(hc, small))
    with pytest.raises(_lib.FcbError):
        cl.query_points("transactions", query=synth.query_vector(spec, 7).tolist(), limit=10000)
    small.close()
