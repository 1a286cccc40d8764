#!/bin/bash
# kernel / e2e time of the bench frame for a few light-grid cell factors (RTB200_GRID_CELL).  For gpurun.
cd "$(dirname "$0")/.."
for f in ${@:-0.25}; do
  RTB200_GRID_CELL=$f python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('cell', '$f', 'ms/frame %.2f' % d['ms_per_step'], 'e2e %.2f' % d['e2e']['ms_per_step'], d['roofline']['per_kernel_ms'], 'build %.2f' % d['roofline']['shadow_queue']['grid_build_ms_per_sample_table'], 'entries', d['roofline']['shadow_queue']['grid_entries'])"
done
