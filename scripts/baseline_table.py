"""BASELINE.md section 3 rows from the committed bench lines (profiles/r2_bench_c<config>_<N>gpu.json)."""
import json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line(c, n):
    p = os.path.join(ROOT, "profiles", "r2_bench_c%s_%dgpu.json" % (c, n))
    return json.loads(open(p).read().strip().splitlines()[-1]) if os.path.exists(p) else None


names = {"3": "(3) big_bunny 4K 16 spp, distance 288", "3b": "(3b) the same, distance 576", "4": "(4) 64 big_bunny copies (317 953 primitives) 4K 1 spp",
         "5": "(5) 120-pose orbit 1080p 1 spp, one step = 120 frames"}
print("| config | device | ms/frame | Mrays/s primary | Mrays/s total | e2e ms | e2e Mrays/s | speed-up vs CPU (device / e2e) | scaling eff. (device / e2e) | roofline.frac |")
print("|---|---|---|---|---|---|---|---|---|---|")
for c in ("3", "3b", "4", "5"):
    one = line(c, 1)
    cpu = one["cpu_baseline"]
    print("| %s | host CPU, %d threads, C++ restatement (\"%s\"): %s | — | — | %.4g | — | — | 1 | — | — |" % (names[c], cpu["cores"], cpu["kind"], cpu["sample"], cpu["value"]))
    for n in (1, 2, 4, 8):
        d = line(c, n)
        if not d:
            continue
        e = d["e2e"]
        eff_d = one["ms_per_step"] / (n * d["ms_per_step"]); eff_e = one["e2e"]["ms_per_step"] / (n * e["ms_per_step"])
        extra = " = **%.0f frames/s**" % d["frames_per_s"] if d.get("frames_per_s") else ""
        print("| (%s) | %d × B200 | %.2f%s | %.0f | %.0f | %.2f | %.0f | %.2g / %.2g | %.2f / %.2f | %.3f |"
              % (c, n, d["ms_per_step"], extra, d.get("mrays_per_s_primary_only", 0), d["value"], e["ms_per_step"], e["value"], d["value"] / cpu["value"], e["value"] / cpu["value"],
                 eff_d, eff_e, d["roofline"]["frac"]))
