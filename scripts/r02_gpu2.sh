#!/bin/bash
# round 2, 1-GPU pass 2: the tests that failed / are new, aggregation A/B, extras
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_general_k.py tests/test_gpu_sharded.py tests/test_gpu_aggregate.py tests/test_gpu_exact.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_pytest2.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest2.log
tail -15 gpurun_out/r02_pytest2.log
rm -f gpurun_out/r02_agg_ab.jsonl
for mode in "default" "FCB_AGG_LARGE=atomic"; do
  echo "== $mode" >> gpurun_out/r02_agg_ab.jsonl
  if [ "$mode" = "default" ]; then timeout 300 python scripts/bench_extra.py --only agg --steps 10 >> gpurun_out/r02_agg_ab.jsonl 2>> gpurun_out/r02_agg_ab.err
  else env $mode timeout 300 python scripts/bench_extra.py --only agg --steps 10 >> gpurun_out/r02_agg_ab.jsonl 2>> gpurun_out/r02_agg_ab.err; fi
done
cut -c1-400 gpurun_out/r02_agg_ab.jsonl
tail -5 gpurun_out/r02_agg_ab.err
