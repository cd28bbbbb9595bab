#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregate.py -m gpu -q > gpurun_out/r02_pytest8.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r02_pytest8.log
rm -f gpurun_out/r02_agg_ab4.jsonl
run() { echo "== $1" >> gpurun_out/r02_agg_ab4.jsonl; env $2 timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | head -1 | cut -c1-300 >> gpurun_out/r02_agg_ab4.jsonl; }
run "in-tree: two cell sets, one barrier per piece" "X=1"
run "one cell set, two barriers per piece" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_onebuf.so"
run "in-tree again" "X=1"
sed 's/"workload".*(\(.*\))", "ms"/"\1", "ms"/' gpurun_out/r02_agg_ab4.jsonl | cut -c1-200
