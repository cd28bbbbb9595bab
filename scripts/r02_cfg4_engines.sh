#!/bin/bash
# configs[3] routed search with 2..4 engines per GPU (experiment): usage r02_cfg4_engines.sh N
cd "$(dirname "$0")/.."
N=${1:-2}
mkdir -p gpurun_out
FCB_CFG4_THREADS=1 FCB_CFG4_ENGINES=4 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 \
  > gpurun_out/r02_cfg4_engines_n$N.json 2> gpurun_out/r02_cfg4_engines_n$N.err; echo "bench rc=$?"
python - <<P
import json
d = json.loads(open("gpurun_out/r02_cfg4_engines_n$N.json").read().strip().splitlines()[-1])
print(json.dumps(d["cfg4"]["tenants_routed"], indent=1)[:1800])
P
tail -5 gpurun_out/r02_cfg4_engines_n$N.err
