#!/bin/bash
# A/B of the LK step under environment switches:  tools/tune_env.sh "AOF_LK_SPLIT=1" "AOF_LK_PARTS=3" ...
for v in "$@"; do
  env $v timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --workload lk 2>/dev/null | python tools/print_kernels.py "[$v]"
done
