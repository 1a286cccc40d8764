#!/bin/bash
# Times the bench frame with every variants/librtb200_*.so (default kernel forms) -- one line per library.  usage: ab_variant_libs.sh [config]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for lib in variants/librtb200_*.so; do
  n=$(basename $lib .so)
  RTB200_LIB=$PWD/$lib timeout 300 python - "$n" "${1:-3}" <<'PY' 2>&1 | tail -1
import os, sys, zlib
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, rtb200, scenes as S
n, cfg = sys.argv[1], sys.argv[2]
v, c = (S.instanced_bunnies(8) if cfg == "4" else S.big_bunny())
kw = dict(width=3840, height=2160, nb_ray=1) if cfg == "4" else dict(width=3840, height=2160, nb_ray=16, distance=576.0 if cfg == "3b" else 288.0)
sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples()); sc.render(**kw)
ov = [sc.render(**kw)[1]["kernel_ms"] for _ in range(4)]
pr = [sc.render(profile=True, **kw) for _ in range(3)]
st = pr[-1][1]
print("%-24s config %s: overlapped %.3f | primary %.3f shadow %.3f shade %.3f | crc %08x" % (n, cfg, min(ov), np.mean([p[1]["primary_ms"] for p in pr]),
      np.mean([p[1]["shadow_ms"] for p in pr]), np.mean([p[1]["shade_ms"] for p in pr]), zlib.crc32(pr[-1][0].tobytes()) & 0xffffffff))
PY
done
