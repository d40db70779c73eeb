#!/bin/bash
# usage: tools/gpu_retry_n.sh <gpus> <log> <timeout_s> <command...>  -- gpurun --gpus N, retried while the pod answers "busy"
n=$1; shift; log=$1; shift; to=$1; shift
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --gpus $n --timeout $to -- "$@" > $log 2>&1
  if ! grep -q "status=transient" $log; then break; fi
  sleep 60
done
