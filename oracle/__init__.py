"""ORACLE — test infrastructure only.

CPU restatements of the reference's hot path used as the parity checker:
  local_mode.py  numpy model of qdrant-client local-mode query_points (PARITY UNPINNED)
  canon.c/.py    plain-C restatement with pinned summation/tie order (bit-exact checker)
  groupby.py     pandas groupby().sum() — the reference's own aggregation library call

Import rule: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import anything from this package. The product
(financial-chatbot-llm_b200/) never does and fails loudly without its CUDA library.
"""
