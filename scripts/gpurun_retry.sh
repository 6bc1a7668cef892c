#!/bin/bash
# usage: gpurun_retry.sh <timeout> '<command>' : gpurun, retried every 2 minutes while the pod answers "transient" (nothing charged)
T=$1; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout $T -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$out"; exit 0
done
echo "gpurun_retry: gave up"; exit 3
