"""Per-kernel split of config 2 (big_bunny 1920x1080 NB_RAY=1) and of one 8-GPU band share of config 3."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S
v, c = S.big_bunny()
sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
for kw in (dict(width=1920, height=1080, nb_ray=1), dict(width=1920, height=1080, nb_ray=4), dict(width=3840, height=2160, nb_ray=1)):
    best = None
    for i in range(4):
        img, st = sc.render(profile=True, **kw)
        if best is None or st["kernel_ms"] < best["kernel_ms"]: best = st
    img, st2 = sc.render(**kw)
    print(kw, {k: round(best[k], 3) for k in ("kernel_ms", "primary_ms", "shadow_ms", "shade_ms")}, "unprofiled %.3f" % st2["kernel_ms"], "queued", best["queued_hits"], "launches", best["launches"])
