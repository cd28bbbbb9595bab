#!/bin/bash
# usage: gpurun_retry.sh <timeout-seconds> [--gpus N] -- <command>   retries while the pod answers "busy" (nothing charged)
TMO=$1; shift
for i in $(seq 1 30); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$TMO" "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 150; continue; fi
  echo "$out" | tail -60
  exit 0
done
echo "gave up after 30 busy answers"
