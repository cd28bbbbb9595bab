#!/bin/bash
# A/B of aggregation builds: in-tree library first (with its tests), then every build_ab/*.so through FCB_LIB_PATH
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregate.py -m gpu -q -x > gpurun_out/r02_agg_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_agg_tests.log
: > gpurun_out/r02_agg_ab.jsonl
for rep in 1; do
  echo "{\"lib\": \"in-tree\"}" >> gpurun_out/r02_agg_ab.jsonl
  timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | cut -c1-400 >> gpurun_out/r02_agg_ab.jsonl
  for so in build_ab/*.so; do
    echo "{\"lib\": \"$so\"}" >> gpurun_out/r02_agg_ab.jsonl
    FCB_LIB_PATH=$PWD/$so timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | cut -c1-400 >> gpurun_out/r02_agg_ab.jsonl
  done
done
cat gpurun_out/r02_agg_ab.jsonl
