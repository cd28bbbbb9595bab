#!/bin/bash
# host-API latency A/B: results written straight into pinned memory (default) vs a device-to-host copy (FCB_HOST_DIRECT=0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_search.py tests/test_gpu_golden.py tests/test_gpu_exact.py tests/test_gpu_config0.py -m gpu -q -x 2>&1 | tail -3
: > gpurun_out/r02_hostpath.jsonl
for d in 1 0 1 0; do
  FCB_HOST_DIRECT=$d timeout 600 python scripts/bench_hostpath.py 2>> gpurun_out/r02_hostpath.err >> gpurun_out/r02_hostpath.jsonl
done
cat gpurun_out/r02_hostpath.jsonl
