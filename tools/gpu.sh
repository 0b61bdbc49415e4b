#!/bin/bash
# [GPUS=2] tools/gpu.sh TIMEOUT 'command' -- gpurun with retries while the pod answers "transient" / busy (nothing is charged for those)
T=$1; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$T" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|status=busy\|retry in a few minutes"; then sleep 90; continue; fi
  echo "$out"; exit 0
done
echo "$out"; exit 3
