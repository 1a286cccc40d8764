#!/bin/bash
# One gpurun call: GPU parity tests + a short bench line without the CPU baseline (A/B of a kernel change).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_quick.json").read().strip().splitlines()[-1])
print("ms/frame %.3f  e2e %.3f  per kernel %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["per_kernel_ms"]))
PY
