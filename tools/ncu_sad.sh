#!/bin/bash
# one ncu --set full capture of the SAD kernel at 4096 and at 32768 pairs per launch (run AFTER the plain command exited 0)
T=${1:-r2}
O=gpurun_out
for n in 4096 32768; do
AOF_BENCH_SAD_STREAMS=$n python bench.py --steps 5 --warmup 3 --no-cpu --workload sad > $O/${T}_sad_plain_$n.log 2>&1 &&
AOF_BENCH_SAD_STREAMS=$n ncu --set full --clock-control none --import-source on -k regex:sad_flow -s 6 -c 2 -o $O/${T}_sad_full_$n -f python bench.py --steps 5 --warmup 3 --no-cpu --workload sad > $O/${T}_sad_ncu_$n.log 2>&1
done
for n in 4096 32768; do
ncu -i $O/${T}_sad_full_$n.ncu-rep --page raw --csv > $O/${T}_sad_full_${n}_raw.csv 2>/dev/null
ncu -i $O/${T}_sad_full_$n.ncu-rep --page source --csv > $O/${T}_sad_full_${n}_source.csv 2>/dev/null
done
ls -la $O | tail
