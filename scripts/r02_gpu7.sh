#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregate.py -m gpu -q > gpurun_out/r02_pytest7.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r02_pytest7.log
rm -f gpurun_out/r02_agg_ab3.jsonl
run() { echo "== $1" >> gpurun_out/r02_agg_ab3.jsonl; env $2 timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | cut -c1-300 >> gpurun_out/r02_agg_ab3.jsonl; }
run "in-tree: lean row path in aggregate_kernel, plain warp-table path" "X=1"
run "branch-free C version in both paths" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_branchfree.so"
run "in-tree again" "X=1"
sed 's/"workload".*(\(.*\))", "ms"/"\1", "ms"/' gpurun_out/r02_agg_ab3.jsonl | cut -c1-200
