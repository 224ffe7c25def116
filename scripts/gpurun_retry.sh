#!/bin/bash
# usage: gpurun_retry.sh TIMEOUT 'command'  [extra gpurun args...] — retries while the pod answers busy / transient
T=$1; CMD=$2; shift 2
for i in $(seq 1 30); do
  OUT=$(/usr/local/graft/bin/gpurun --timeout "$T" "$@" -- "$CMD" 2>&1)
  if echo "$OUT" | grep -q "status=transient\|status=busy\|rc=3\b"; then sleep 90; continue; fi
  echo "$OUT"; exit 0
done
echo "$OUT"; exit 3
