#!/bin/bash
# usage: tools/tune_env.sh VAR "v1 v2 ..." [bench args]  -- runs bench.py once per value of VAR and prints the
# per-kernel event times (tuning aid, not a bench result)
var=$1; vals=$2; shift 2
for v in $vals; do
  env $var=$v python bench.py --no-cpu --steps 30 --warmup 3 "$@" 2>/dev/null | python tools/print_kernels.py "$var=$v"
done
