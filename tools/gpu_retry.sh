#!/bin/bash
# usage: tools/gpu_retry.sh <log> <timeout_s> <command...>   -- retries while the pod answers "busy" (nothing is charged)
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1
  if ! grep -q "status=transient" $log; then break; fi
  sleep 45
done
