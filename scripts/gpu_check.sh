#!/bin/bash
# One gpurun call: the whole GPU test suite, the smoke test, and the bench line of a config.  usage: gpu_check.sh [config=3] [steps=8]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
C=${1:-3}
timeout 900 python bench.py --config $C --steps ${2:-8} > gpurun_out/bench_c$C.json 2> gpurun_out/bench_c$C.err; echo "bench rc=$?"
python - gpurun_out/bench_c$C.json <<'PY'
import json, sys
j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r=j["roofline"]
print("value %.0f ms %.3f e2e %.0f e2e_ms %.3f frac %.3f cpu %s" % (j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["ms_per_step"], r["frac"], j.get("cpu_baseline", {}).get("value")))
print({k: (round(v["ms"], 3), round(v["frac"], 3)) for k, v in r.get("per_kernel", {}).items()}, {k: j[k] for k in j if "match" in k or "redo" in k or "frames_per" in k})
PY
