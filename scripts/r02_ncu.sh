#!/bin/bash
# round 2 ncu pass (1 GPU): launch list of the bench command, full capture of the scan kernel, full capture of the
# shared-atomic aggregation kernel (its atom metrics), per the profiling recipe (plain run first, same command line).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --headline-only --no-cpu-baseline"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_steps5.csv $CMD > gpurun_out/r02_ncu_a.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:scan_topk_kernel -s 50 -c 1 -o gpurun_out/r02_scan_topk $CMD > gpurun_out/r02_ncu_b.log 2>&1
echo "scan capture rc=$?"
AGG="python scripts/bench_extra.py --only agg --steps 2"
$AGG > gpurun_out/r02_ncu_plain_agg.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:aggregate_kernel -s 1 -c 1 -o gpurun_out/r02_aggregate_atomic $AGG > gpurun_out/r02_ncu_c.log 2>&1
echo "agg capture rc=$?"
ls -la gpurun_out/*.ncu-rep
tail -3 gpurun_out/r02_ncu_a.log gpurun_out/r02_ncu_b.log gpurun_out/r02_ncu_c.log
