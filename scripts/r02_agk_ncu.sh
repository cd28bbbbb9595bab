#!/bin/bash
# one full ncu capture of the warp-table aggregation kernel in its date form (FCB_AGG_KEYS=0: the shard gets no keys, so every
# launch of the bench is the date form), after the same command has run plainly; then the default bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
AGG="python scripts/bench_extra.py --only agg --steps 2"
FCB_AGG_KEYS=0 $AGG > gpurun_out/r02_ncu_plain_agd.log 2>&1 &&
FCB_AGG_KEYS=0 ncu --set full --clock-control none --import-source on -k regex:aggregate_warp_kernel -s 1 -c 1 -f -o gpurun_out/r02_aggregate_warp_date $AGG > gpurun_out/r02_ncu_agd.log 2>&1
echo "date-form capture rc=$?"; tail -2 gpurun_out/r02_ncu_agd.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc=$?"
python - <<'P'
import json
d = json.loads(open("gpurun_out/r02_bench.json").read().strip().splitlines()[-1])
x = d["extra"][1]
print(d["value"], d["e2e"]["value"], x["ms"], x["roofline"]["frac"], x["without_group_keys"], x["parity"], d["extra"][0]["ms"])
P
