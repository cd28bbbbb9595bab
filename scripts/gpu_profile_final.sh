#!/bin/bash
# Round-end evidence, one profiler pass per invocation:
#   gpu_profile_final.sh tests     full GPU tests, headline bench (+ reference arm), secondary workloads
#   gpu_profile_final.sh launches  launch list of the headline bench command (after a plain run of the same command)
#   gpu_profile_final.sh scan      ncu --set full of the scan kernel for the same command
#   gpu_profile_final.sh dense     ncu --set full of the batched tensor-core kernel (scripts/run_one.py cfg3)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
case "${1:-tests}" in
tests)
    python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/final_pytest.log
    python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench exit $?"
    python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/final_bench_reference.json 2>/dev/null; echo "ref exit $?"
    python scripts/bench_extra.py > gpurun_out/final_extra.jsonl 2>/dev/null; echo "extra exit $?"
    ;;
launches)
    $CMD > gpurun_out/final_plain.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/final_launches.csv $CMD > /dev/null 2>&1
    echo "launch list exit $?"
    ;;
scan)
    $CMD > gpurun_out/final_plain2.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:scan_topk -s 3 -c 2 -o gpurun_out/final_prof_scan $CMD > /dev/null 2>&1
    echo "full exit $?"
    ;;
dense)
    python scripts/run_one.py cfg3 3 > gpurun_out/final_plain3.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:dense_topk2 -s 1 -c 1 -o gpurun_out/final_prof_dense python scripts/run_one.py cfg3 3 > /dev/null 2>&1
    echo "dense full exit $?"
    ;;
esac
