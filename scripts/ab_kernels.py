"""Kernel times of a bench frame (gpurun): per-kernel times (RTB_FLAG_PROFILE, chunks in sequence) and the overlapped frame
time, with a CRC of the frame, for a set of scene options (A/B of tunables).
usage: python scripts/ab_kernels.py [config: 3|3b|4] [extra option=value ...]"""
import os, sys, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rtb200, scenes as S

cfg = sys.argv[1] if len(sys.argv) > 1 else "3"
extra = {}
for kv in sys.argv[2:]:
    k, v = kv.split("=")
    extra[k] = float(v) if "." in v else int(v)
if cfg == "4":
    v, c = S.instanced_bunnies(8); kw = dict(width=3840, height=2160, nb_ray=1)
else:
    v, c = S.big_bunny(); kw = dict(width=3840, height=2160, nb_ray=16, distance=576.0 if cfg == "3b" else 288.0)
smp = S.make_samples()
crcs = set()
for _ in range(1):
    sc = rtb200.Scene(v, c, **extra)
    sc.set_samples(smp)
    sc.render(**kw)
    ov, pr, sh, sd = [], [], [], []
    for _ in range(4):
        img, st = sc.render(**kw)
        ov.append(st["kernel_ms"])
    for _ in range(3):
        img, st = sc.render(profile=True, **kw)
        pr.append(st["primary_ms"]); sh.append(st["shadow_ms"]); sd.append(st["shade_ms"])
    crc = zlib.crc32(img.tobytes()) & 0xffffffff
    crcs.add(crc)
    print("config %s %s: overlapped %.3f ms | primary %.3f shadow %.3f shade %.3f | crc %08x queued %d grid_build %.2f ms"
          % (cfg, extra, min(ov), np.mean(pr), np.mean(sh), np.mean(sd), crc, st["queued_hits"], st["grid_build_ms"]), flush=True)
    del sc
