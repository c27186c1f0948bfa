#!/bin/bash
# usage: tools/gpu/retry.sh <timeout-seconds> <gpus> '<command run on the GPU box>'
# Calls gpurun until it is not answered "transient" (busy pod: nothing charged).
T=$1; G=$2; CMD=$3
for i in $(seq 1 40); do
  if [ "$G" = "1" ]; then OUT=$(/usr/local/graft/bin/gpurun --timeout $T -- "$CMD" 2>&1); else OUT=$(/usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$CMD" 2>&1); fi
  if echo "$OUT" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$OUT" | tail -60
  exit 0
done
echo "gave up after 40 transient answers"
