#!/bin/bash
# lazy-FAST statistics, then one ncu --set full capture of the LK kernels of a steady-state step (run AFTER the plain run exited 0)
T=${1:-r2}
K=${2:-lk_fast_pick}
O=gpurun_out
python tools/lazy_probe.py 512 8 > $O/${T}_lazy_stats.txt 2>&1
tail -9 $O/${T}_lazy_stats.txt
python bench.py --steps 5 --warmup 3 --no-cpu --workload lk > $O/${T}_lk_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 1 -o $O/${T}_lk_full -f python bench.py --steps 5 --warmup 3 --no-cpu --workload lk > $O/${T}_lk_ncu.log 2>&1
ncu -i $O/${T}_lk_full.ncu-rep --page raw --csv > $O/${T}_lk_full_raw.csv 2>/dev/null
ncu -i $O/${T}_lk_full.ncu-rep --page source --csv > $O/${T}_lk_full_source.csv 2>/dev/null
ls -la $O | tail -5
