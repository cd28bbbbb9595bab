#!/bin/bash
# A/B of the group-key aggregation builds, then one full ncu capture of aggregate_key_kernel (plain run first, same command)
cd "$(dirname "$0")/.."
bash scripts/r02_agg_ab.sh
AGG="python scripts/bench_extra.py --only agg --steps 2"
$AGG > gpurun_out/r02_ncu_plain_agk.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:aggregate_key_kernel -s 1 -c 1 -f -o gpurun_out/r02_aggregate_key $AGG > gpurun_out/r02_ncu_agk.log 2>&1
echo "agk capture rc=$?"; tail -2 gpurun_out/r02_ncu_agk.log
