#!/bin/bash
# Round-2 evidence in one gpurun call: bench lines of every config on 1 GPU, ncu --set full of the three hot kernels (config 3;
# k_shadow_grid also on 3b and 4), and the ncu launch list of the bench -- each ncu run after the same command exited 0 without it.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for C in 3 3b 4 5; do
  timeout 900 python bench.py --config $C --steps 8 > gpurun_out/r2_bench_c${C}_1gpu.json 2> gpurun_out/r2_bench_c${C}_1gpu.err; echo "bench $C rc=$?"
done
bash scripts/gpu_ncu_r2.sh
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
python - <<'PY'
import json
for C in ("3", "3b", "4", "5"):
    try:
        j = json.loads(open("gpurun_out/r2_bench_c%s_1gpu.json" % C).read().strip().splitlines()[-1]); r = j["roofline"]
        print(C, "value %.0f ms %.3f e2e_ms %.3f frac %.3f cpu %s" % (j["value"], j["ms_per_step"], j["e2e"]["ms_per_step"], r["frac"], j.get("cpu_baseline", {}).get("value")),
              {k: (round(v["ms"], 3), round(v["frac"], 3)) for k, v in r.get("per_kernel", {}).items()}, {k: j[k] for k in j if "match" in k or "frames_per" in k})
    except Exception as e:
        print(C, "no line", e)
PY
