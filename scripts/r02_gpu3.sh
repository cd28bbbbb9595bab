#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_aggregate.py tests/test_serving.py -m gpu -q > gpurun_out/r02_pytest3.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest3.log
tail -8 gpurun_out/r02_pytest3.log
rm -f gpurun_out/r02_agg_ab.jsonl
for mode in claim bucket atomic; do
  echo "== FCB_AGG_LARGE=$mode" >> gpurun_out/r02_agg_ab.jsonl
  FCB_AGG_LARGE=$mode timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | head -1 | cut -c1-330 >> gpurun_out/r02_agg_ab.jsonl
done
cat gpurun_out/r02_agg_ab.jsonl; tail -5 gpurun_out/r02_agg_ab.err
timeout 600 python scripts/bench_serving.py --queries 1500 > gpurun_out/r02_serving.jsonl 2> gpurun_out/r02_serving.err; echo "serving rc=$?"; cat gpurun_out/r02_serving.jsonl; tail -5 gpurun_out/r02_serving.err
