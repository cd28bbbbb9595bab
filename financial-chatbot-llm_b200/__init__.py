"""B200-native retrieval + aggregation engine for the transaction lookup of
kyshu11027/financial-chatbot-llm (tools/qdrant_tool.py, tools/plot_tool.py).

Import as ``financial_chatbot_llm_b200``. Submodules are imported lazily so that
host-only pieces (synth, qdrant_models) work without the CUDA library; anything
that computes goes through the C-ABI library and raises if it is missing.
"""
__version__ = "0.1.0"
