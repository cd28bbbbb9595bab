#!/bin/bash
# round 2, multi-GPU pass: usage r02_multi.sh N [tests]
cd "$(dirname "$0")/.."
N=${1:-2}
mkdir -p gpurun_out
if [ "$2" = "tests" ]; then
  timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q > gpurun_out/r02_pytest_multi_n$N.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest_multi_n$N.log
  tail -12 gpurun_out/r02_pytest_multi_n$N.log
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 \
  > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "bench rc=$?"
python - <<P
import json
try:
    d = json.loads(open("gpurun_out/r02_bench_n$N.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "parity_spot_check", "gpu_launches")}, d["roofline"]["frac"], d["roofline"]["kernel_ms"])
    print(json.dumps(d.get("cfg4"), indent=1)[:2500])
except Exception as e:
    print("bench parse failed", e)
print(open("gpurun_out/r02_bench_n$N.err").read()[-2500:])
P
