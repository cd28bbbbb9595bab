#!/bin/bash
# aggregation tuning: default (shared cells) vs everything through global 64-bit reductions
python -m pytest tests/test_gpu_aggregate.py -m gpu -x -q 2>&1 | tail -3
python scripts/bench_extra.py --only agg 2>&1 | grep cfg5
FCB_AGG_SMALL_PIECE=1000000000 python scripts/bench_extra.py --only agg 2>&1 | grep cfg5
