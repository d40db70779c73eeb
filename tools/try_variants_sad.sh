#!/bin/bash
# tuning aid: run the SAD bench (incl. the dense leg) once per prebuilt library variant in tmp_variants/
cp aero-optical-flow_b200/libaof_b200.so /tmp/keep.so
for f in tmp_variants/lib_*.so; do
  cp $f aero-optical-flow_b200/libaof_b200.so
  python bench.py --no-cpu --steps 300 --warmup 3 --workload sad 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$(basename $f)', 'sad_kernel_us=%.2f' % (d['kernels']['sad_flow_kernel']['ms_per_step']*1e3), 'value=%.1fM' % (d['value']/1e6), 'dense_pairs_per_s=%.0f' % d['dense']['pairs_per_s'], 'dense_frac=%.3f' % d['dense']['roofline_int_alu']['frac'])"
done
cp /tmp/keep.so aero-optical-flow_b200/libaof_b200.so
