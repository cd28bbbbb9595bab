import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


@pytest.fixture(scope="session")
def built():
    """Make sure libfcb200.so and liboracle are built (cross-compiles without a GPU)."""
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def small_corpus():
    from financial_chatbot_llm_b200 import synth
    spec = synth.SynthSpec(n_rows=6000, dim=256, n_users=40, seed=11)
    return synth.make_host_corpus(spec)
