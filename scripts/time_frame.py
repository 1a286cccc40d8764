"""GPU timing aid: per-kernel ms of the 4K/16spp bench frame (and 1080p/1spp) for the library in RTB200_LIB, plus a
quick oracle parity check on a pixel subset.  usage: python scripts/time_frame.py [nocheck]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rtb200, scenes as S

v, c = S.big_bunny()
smp = S.make_samples()
sc = rtb200.Scene(v, c); sc.set_samples(smp)
tag = os.path.basename(os.environ.get("RTB200_LIB", "default"))
if "nocheck" not in sys.argv:
    from oracle import oracle as O
    pix = np.concatenate([S.stratified_pixels(1920, 1080, 30, seed=5),
                          [[x, y] for x in range(820, 960, 7) for y in range(440, 700, 7)]]).astype(np.uint32)
    g = sc.render_pixels(pix)
    o = O.OracleScene(v, c, True).render(1920, 1080, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
    bad = int((g["rgb_lin"].view(np.uint32) != o["rgb_lin"].view(np.uint32)).sum() + (g["prim0"] != o["prim0"]).sum())
    print(tag, "parity mismatches:", bad, "of", len(pix), flush=True)
for (w, h, nr) in ((1920, 1080, 1), (3840, 2160, 16)):
    best = None; fast = 1e9
    for it in range(4):
        img, st = sc.render(width=w, height=h, nb_ray=nr, profile=True)
        if best is None or st["kernel_ms"] < best["kernel_ms"]: best = st
        img, st = sc.render(width=w, height=h, nb_ray=nr)
        fast = min(fast, st["kernel_ms"])
    print("   grid: cells %d entries %d big %d build %.2f ms" % (best["grid_cells"], best["grid_entries"], best["grid_big"], best["grid_build_ms"]))
    print("%s %dx%d spp%d: overlapped %.2f ms | serial %.2f ms  primary %.2f  shadow %.2f  shade %.2f  queued %d hits -> %d sub-items (%d walks) of %d hits" % (
        tag, w, h, nr, fast, best["kernel_ms"], best["primary_ms"], best["shadow_ms"], best["shade_ms"], best["queued_hits"], best["queued_groups"], best["queued_walks"], best["primary_hits"]), flush=True)
