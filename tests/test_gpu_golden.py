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
