#!/bin/bash
# tuning aid: LK parity subset + LK bench once per prebuilt library variant in tmp_variants/
cp aero-optical-flow_b200/libaof_b200.so /tmp/keep.so
for f in tmp_variants/lib_*.so; do
  cp $f aero-optical-flow_b200/libaof_b200.so
  timeout 300 python -m pytest tests/test_lk_gpu.py -x -q 2>&1 | tail -1
  timeout 300 python bench.py --no-cpu --steps 50 --warmup 3 --workload lk 2>/dev/null | python tools/print_kernels.py $(basename $f) | cut -c1-330
done
cp /tmp/keep.so aero-optical-flow_b200/libaof_b200.so
