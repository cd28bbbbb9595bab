#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/r02_agg_ab5.jsonl
run() { echo "== $1" >> gpurun_out/r02_agg_ab5.jsonl; env $2 timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | head -1 | cut -c1-300 >> gpurun_out/r02_agg_ab5.jsonl; }
run "in-tree (4 CTAs per SM)" "X=1"
run "5 CTAs per SM" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_ctas5.so"
run "6 CTAs per SM" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_ctas6.so"
run "8 CTAs per SM" "FCB_LIB_PATH=$PWD/build_ab/libfcb200_ctas8.so"
run "in-tree again" "X=1"
sed 's/"workload".*(\(.*\))", "ms"/"\1", "ms"/' gpurun_out/r02_agg_ab5.jsonl | cut -c1-160
AGG="python scripts/bench_extra.py --only agg --steps 2"
$AGG > gpurun_out/r02_ncu_plain_agg.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:aggregate_kernel -s 1 -c 1 -o gpurun_out/r02_aggregate_atomic_v2 $AGG > gpurun_out/r02_ncu_c2.log 2>&1
echo "agg capture rc=$?"
