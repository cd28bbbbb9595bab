#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregate.py tests/test_gpu_sharded.py -m gpu -q > gpurun_out/r02_pytest4.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02_pytest4.log
rm -f gpurun_out/r02_agg_block_ab.jsonl
for lib in in-tree build_ab/libfcb200_agg_1_2.so build_ab/libfcb200_agg_16_4.so build_ab/libfcb200_agg_64_2.so in-tree; do
  echo "== $lib" >> gpurun_out/r02_agg_block_ab.jsonl
  if [ "$lib" = "in-tree" ]; then timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | cut -c1-300 >> gpurun_out/r02_agg_block_ab.jsonl
  else FCB_LIB_PATH=$PWD/$lib timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | cut -c1-300 >> gpurun_out/r02_agg_block_ab.jsonl; fi
done
cat gpurun_out/r02_agg_block_ab.jsonl; tail -3 gpurun_out/r02_agg_ab.err
