#!/bin/bash
# tuning aid: dense SAD parity + bench per prebuilt library variant in tmp_variants/
cp aero-optical-flow_b200/libaof_b200.so /tmp/keep.so
for f in tmp_variants/lib_*.so; do
  cp $f aero-optical-flow_b200/libaof_b200.so
  timeout 300 python -m pytest tests/test_dense.py -x -q -m gpu 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --steps 20 --warmup 3 --workload sad 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l)['dense']; print('$(basename $f)', 'pairs/s %.0f' % d['pairs_per_s'], 'frac %.3f' % d['roofline']['frac'])"
done
cp /tmp/keep.so aero-optical-flow_b200/libaof_b200.so
