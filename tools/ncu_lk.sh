#!/bin/bash
# FAST statistics, then one ncu --set full capture of an LK kernel of a steady-state step (run AFTER the plain run exited 0)
#   tools/ncu_lk.sh <tag> <kernel regex>
T=${1:-r2}
K=${2:-lk_fast}
O=gpurun_out
timeout 200 python tools/lazy_probe.py 512 8 > $O/${T}_lazy_stats.txt 2>&1
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --workload lk > $O/${T}_lk_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 1 -o $O/${T}_lk_full -f python bench.py --steps 5 --warmup 3 --no-cpu --workload lk > $O/${T}_lk_ncu.log 2>&1
ls -la $O | tail -5
