#!/bin/bash
# Round-end evidence: full GPU tests, headline bench, launch list + full ncu capture of the scan kernel for the SAME command.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/final_pytest.log
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/final_bench_reference.json 2>/dev/null; echo "ref exit $?"
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/final_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/final_launches.csv $CMD > /dev/null 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/final_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:scan_topk -s 3 -c 2 -o gpurun_out/final_prof_scan $CMD > /dev/null 2>&1
echo "full exit $?"
python scripts/bench_extra.py > gpurun_out/final_extra.jsonl 2>/dev/null; echo "extra exit $?"
