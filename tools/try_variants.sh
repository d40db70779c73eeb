#!/bin/bash
# tuning aid: run the LK bench once per prebuilt library variant in tmp_variants/
cp aero-optical-flow_b200/libaof_b200.so /tmp/keep.so
for f in tmp_variants/lib_*.so; do
  cp $f aero-optical-flow_b200/libaof_b200.so
  python bench.py --no-cpu --steps 50 --warmup 3 --workload lk 2>/dev/null | python tools/print_kernels.py $(basename $f)
done
cp /tmp/keep.so aero-optical-flow_b200/libaof_b200.so
