#!/bin/bash
# SAD parity tests + SAD bench at 4096 and 32768 pairs per launch
timeout 600 python -m pytest tests/test_sad_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
for n in 4096 32768; do
AOF_BENCH_SAD_STREAMS=$n timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu --workload sad 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']; print('n=$n', 'kernel_us=%.2f' % (r['kernel_ms'] * 1e3), 'value=%.1fM' % (d['value'] / 1e6), 'alu_frac=%.3f' % r['frac'], 'step_us=%.2f' % (d['ms_per_step'] * 1e3))"
done
