#!/bin/bash
# tuning aid: parity + SAD bench once per prebuilt library variant in tmp_variants/ at 4096 and 32768 pairs per launch
cp aero-optical-flow_b200/libaof_b200.so /tmp/keep.so
for f in tmp_variants/lib_*.so; do
  cp $f aero-optical-flow_b200/libaof_b200.so
  timeout 300 python -m pytest tests/test_sad_gpu.py -x -q 2>&1 | tail -1
  for n in 4096 32768; do
  AOF_BENCH_SAD_STREAMS=$n timeout 300 python bench.py --no-cpu --steps 300 --warmup 3 --workload sad 2>>gpurun_out/variants.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('$(basename $f)', 'n=$n', 'sad_kernel_us=%.2f' % (r['kernel_ms']*1e3), 'value=%.1fM' % (d['value']/1e6), 'alu_frac=%.3f' % r['frac'])"
  done
done
cp /tmp/keep.so aero-optical-flow_b200/libaof_b200.so
