#!/bin/bash
# pyramid kernel tuning sweep: occupancy bound x bands (tuning aid)
python -m pytest tests/test_lk_gpu.py -x -q -m gpu -k pyramid 2>&1 | tail -3
for o in 16 20 24; do for b in 1 2 3 4; do
  AOF_PYR_OCC=$o AOF_PYR_BANDS=$b python bench.py --no-cpu --steps 30 --warmup 3 --workload lk 2>/dev/null | python tools/print_kernels.py "occ=$o bands=$b"
done; done
