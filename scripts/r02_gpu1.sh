#!/bin/bash
# round 2, 1-GPU pass: parity tests, headline bench (+extras), cfg3 certificate A/B
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest.log
tail -15 gpurun_out/r02_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc=$?"
rm -f gpurun_out/r02_cfg3_ab.jsonl
for mode in "default" "FCB_DENSE_RIGOROUS=0"; do
  echo "== $mode" >> gpurun_out/r02_cfg3_ab.jsonl
  if [ "$mode" = "default" ]; then timeout 300 python scripts/bench_cfg3.py 30 >> gpurun_out/r02_cfg3_ab.jsonl 2>> gpurun_out/r02_cfg3_ab.err
  else env $mode timeout 300 python scripts/bench_cfg3.py 30 >> gpurun_out/r02_cfg3_ab.jsonl 2>> gpurun_out/r02_cfg3_ab.err; fi
done
cat gpurun_out/r02_cfg3_ab.jsonl
python - <<'P'
import json
try:
    d = json.loads(open("gpurun_out/r02_bench.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "parity_spot_check", "gpu_launches")}, d["roofline"]["frac"], d["e2e"])
    for x in d.get("extra", []):
        print(x.get("workload", "")[:40], x.get("ms"), x.get("roofline", {}).get("frac"), x.get("parity"), x.get("error"), x.get("uncertified_rerun_by_scan"))
    print(d.get("cpu_baseline", {}).get("sample"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/r02_bench.err").read()[-3000:])
P
