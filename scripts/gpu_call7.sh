#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AB_PAIRS="2,2"
( python scripts/ab_kernels.py 3; RTB200_NO_COVER_ORDER=1 python scripts/ab_kernels.py 3; python scripts/ab_kernels.py 3 grid_cell_factor=0.18;  python scripts/ab_kernels.py 3 grid_cell_factor=0.15; python scripts/ab_kernels.py 4 ) 2>&1 | grep -v identical > gpurun_out/ab_span.txt; cat gpurun_out/ab_span.txt
timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout 600 python bench.py --steps 5 --no-cpu-baseline > gpurun_out/bench_c3b.json 2> gpurun_out/bench_c3.err; python - <<'PY'
import json
j=json.loads(open("gpurun_out/bench_c3b.json").read().strip().splitlines()[-1])
print("value %.0f ms %.3f e2e_ms %.3f frac %.3f build %.2f" % (j["value"], j["ms_per_step"], j["e2e"]["ms_per_step"], j["roofline"]["frac"], j["roofline"]["shadow_queue"]["grid_build_ms_per_sample_table"]), j["roofline"]["shadow_queue"])
PY
