#!/bin/bash
# One GPU-box pass that regenerates everything profiles/ is built from.  usage: tools/full_refresh.sh <tag>
# Every ncu pass runs only after the same command has exited 0 without ncu.
T=${1:-rX}; O=gpurun_out
nvidia-smi -L; nproc
python -m pytest tests -m gpu -x -q > $O/${T}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${T}_pytest_gpu.log; tail -3 $O/${T}_pytest_gpu.log
python __graft_entry__.py smoke > $O/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/${T}_smoke.log
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_ref.json 2>&1; echo "ref exit $?"
python bench.py --steps 3 --warmup 3 --no-cpu > $O/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > $O/${T}_ncu1.log 2>&1
# full capture of two steady-state LK steps (7 launches each: pyramid, track, plan, masked band, select, flagged band, select)
python bench.py --steps 3 --warmup 3 --no-cpu --workload lk > $O/${T}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:lk_ -s 21 -c 14 -o $O/${T}_lk_full -f python bench.py --steps 3 --warmup 3 --no-cpu --workload lk > $O/${T}_ncu2.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu --workload sad > $O/${T}_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sad_flow -s 5 -c 2 -o $O/${T}_sad_full -f python bench.py --steps 3 --warmup 3 --no-cpu --workload sad > $O/${T}_ncu3.log 2>&1
ls -la $O | grep ${T}_
