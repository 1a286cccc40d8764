"""k_primary with 1 / 2 / 4 warps per 32-pixel tile (RTB200_PRIMARY_SPLIT, default: by chunk size): kernel times of rank 0's band share of the
4K / 16 spp frame for world = 1, 2, 4, 8 on one GPU."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S
v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
for world in (8, 4, 2, 1):
    kw = dict(width=3840, height=2160, nb_ray=16)
    if world > 1: kw.update(band_rows=4, rank=0, world=world)
    for _ in range(4): sc.render(**kw)
    best = min(sc.render(**kw)[1]["kernel_ms"] for _ in range(6))
    st = sc.render(profile=True, **kw)[1]
    print("split %s world %d: frame %.3f ms | primary %.3f shadow %.3f shade %.3f" % (os.environ.get("RTB200_PRIMARY_SPLIT", "auto"), world, best, st["primary_ms"], st["shadow_ms"], st["shade_ms"]), flush=True)
