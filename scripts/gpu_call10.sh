#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export AB_PAIRS="2,2"
( python scripts/ab_kernels.py 3; python scripts/ab_kernels.py 3 grid_cell_factor=0.18; python scripts/ab_kernels.py 4 ) 2>&1 | grep -v identical
timeout 1800 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for env in "" "RTB200_NO_FILL_OVERLAP=1" "RTB200_GRID_CELL=0.18"; do
env $env timeout 600 python bench.py --steps 8 --no-cpu-baseline > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp.err; python - "$env" <<'PY'
import json, sys
j=json.loads(open("gpurun_out/bench_tmp.json").read().strip().splitlines()[-1])
print(sys.argv[1], "value %.0f ms %.3f e2e_ms %.3f frac %.3f build %.2f" % (j["value"], j["ms_per_step"], j["e2e"]["ms_per_step"], j["roofline"]["frac"], j["roofline"]["shadow_queue"]["grid_build_ms_per_sample_table"]), j.get("e2e_host_frame_matches_single_gpu"), j.get("e2e_grid_redo_last_step"))
PY
done
