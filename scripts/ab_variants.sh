#!/bin/bash
# Per-rank (world 8, ranks averaged) and full-frame kernel times for every variant library under variants/ (gpurun).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for lib in variants/librtb200_*.so; do
  n=$(basename $lib .so)
  RTB200_SHADE_PHASES=1 RTB200_LIB=$PWD/$lib python scripts/band_times.py 8 16777216 > gpurun_out/var_w8_$n.log 2>&1
  RTB200_SHADE_PHASES=1 RTB200_LIB=$PWD/$lib python scripts/band_times.py 1 16777216 > gpurun_out/var_w1_$n.log 2>&1
  python - $n <<'PY'
import re, sys
n = sys.argv[1]
for w in ("8", "1"):
    rows = [l for l in open("gpurun_out/var_w%s_%s.log" % (w, n)) if "overlapped" in l]
    if not rows:
        print(n, "world", w, "no rows"); continue
    f = lambda k: [float(re.search(k + r" ([\d.]+)", l).group(1)) for l in rows]
    ov, pr, sh, sd = f("overlapped"), f("primary"), f("shadow"), f("shade")
    m = lambda v: sum(v) / len(v)
    print("%-28s world %s: overlapped %.3f | primary %.3f shadow %.3f shade %.3f" % (n, w, m(ov), m(pr), m(sh), m(sd)))
PY
done
