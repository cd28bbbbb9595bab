#!/bin/bash
# round 2, first GPU pass: parity tests, headline bench, cfg3 certificate A/B
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest1.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest1.log
tail -5 gpurun_out/r02_pytest1.log
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/r02_bench1.json 2> gpurun_out/r02_bench1.err; echo "bench rc=$?"
for mode in "default" "FCB_DENSE_QFMT=bf16" "FCB_DENSE_RIGOROUS=0" "FCB_DENSE_QFMT=bf16 FCB_DENSE_RIGOROUS=0"; do
  echo "== $mode" >> gpurun_out/r02_cfg3_ab.jsonl
  if [ "$mode" = "default" ]; then timeout 300 python scripts/bench_cfg3.py 30 >> gpurun_out/r02_cfg3_ab.jsonl 2>> gpurun_out/r02_cfg3_ab.err
  else env $mode timeout 300 python scripts/bench_cfg3.py 30 >> gpurun_out/r02_cfg3_ab.jsonl 2>> gpurun_out/r02_cfg3_ab.err; fi
done
cat gpurun_out/r02_cfg3_ab.jsonl
head -c 1500 gpurun_out/r02_bench1.json
