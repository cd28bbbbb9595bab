#!/bin/bash
# end of round 2: launch list of the headline bench command and one full capture of the scan kernel (plain run first, same command)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --headline-only --no-cpu-baseline"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_steps5.csv $CMD > gpurun_out/r02_ncu_a.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:scan_topk_kernel -s 50 -c 1 -f -o gpurun_out/r02_scan_topk $CMD > gpurun_out/r02_ncu_b.log 2>&1
echo "scan capture rc=$?"
tail -n 2 gpurun_out/r02_ncu_a.log; tail -n 2 gpurun_out/r02_ncu_b.log
