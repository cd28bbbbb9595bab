#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_full.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest_full.log
tail -6 gpurun_out/r02_pytest_full.log
timeout 300 python scripts/bench_extra.py --only agg --steps 10 2>> gpurun_out/r02_agg_ab.err | cut -c1-300 > gpurun_out/r02_agg_final.jsonl; cat gpurun_out/r02_agg_final.jsonl
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc=$?"
python - <<'P'
import json
try:
    d = json.loads(open("gpurun_out/r02_bench.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "parity_spot_check", "gpu_launches")}, d["roofline"]["frac"], d["roofline"]["traffic"])
    for x in d.get("extra", []):
        print(x.get("workload", "")[:40], x.get("ms"), x.get("roofline", {}).get("frac"), x.get("parity"), x.get("error"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/r02_bench.err").read()[-3000:])
P
python -c "import __graft_entry__ as g; g.smoke()"
