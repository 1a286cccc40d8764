"""Quick GPU sanity run (development aid): parity on a pixel subset + timings."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rtb200, scenes as S
from oracle import oracle as O

v, c = S.big_bunny()
smp = S.make_samples()
t = time.time(); sc = rtb200.Scene(v, c); print("scene_create s", time.time() - t, sc.info(), flush=True)
sc.set_samples(smp)
pix = S.stratified_pixels(1920, 1080, 24)
g = sc.render_pixels(pix, counters=True)
print("gpu list stats", g["stats"], flush=True)
osc = O.OracleScene(v, c, True)
o = osc.render(1920, 1080, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
print("oracle stats", o["stats"], flush=True)
for k in ("hit_count", "prim0", "rgb8"):
    print(k, "mismatch", int((g[k] != o[k]).sum()), "of", g[k].size)
print("t0 bit mismatch", int((g["t0"].view(np.uint32) != o["t0"].view(np.uint32)).sum()))
print("rgb_lin bit mismatch", int((g["rgb_lin"].view(np.uint32) != o["rgb_lin"].view(np.uint32)).sum()),
      "max abs", float(np.abs(g["rgb_lin"] - o["rgb_lin"]).max()))
for (w, h, nr) in ((1920, 1080, 1), (3840, 2160, 1), (3840, 2160, 16)):
    for it in range(3):
        img, st = sc.render(width=w, height=h, nb_ray=nr)
    print(w, h, nr, st, flush=True)
    rays = st["primary_rays"] + st["shadow_rays"]
    print("   Mrays/s total", rays / st["kernel_ms"] / 1e3)
img, st = sc.render(width=1920, height=1080, counters=True)
print("counters", st)
rtb200.write_png(os.path.join(ROOT, "gpurun_out", "big_bunny_1080p.png"), img)
