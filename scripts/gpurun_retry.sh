#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit 3 / status=transient; nothing is charged).
#   scripts/gpurun_retry.sh <timeout-seconds> '<command>' [--gpus N]
T=$1; CMD=$2; shift 2
for attempt in $(seq 1 30); do
  OUT=$(/usr/local/graft/bin/gpurun "$@" --timeout $T -- "$CMD" 2>&1)
  if echo "$OUT" | grep -q "status=transient"; then sleep 150; continue; fi
  echo "$OUT"; exit 0
done
echo "gave up after 30 attempts"; exit 3
