#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregate.py -m gpu -q > gpurun_out/r02_pytest5.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r02_pytest5.log
rm -f gpurun_out/r02_agg_ab2.jsonl
run() { echo "== $1" >> gpurun_out/r02_agg_ab2.jsonl; env $2 timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | head -1 | cut -c1-300 >> gpurun_out/r02_agg_ab2.jsonl; }
run "in-tree (concurrent launches, unconditional presence store)" "X=1"
run "in-tree, FCB_AGG_CONCURRENT=0" "FCB_AGG_CONCURRENT=0"
run "presence check restored" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_presentcheck.so"
run "all users through the warp-table path" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_allwarp.so"
run "in-tree again" "X=1"
cat gpurun_out/r02_agg_ab2.jsonl; tail -3 gpurun_out/r02_agg_ab.err
