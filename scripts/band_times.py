"""Kernel time of ONE rank's share of the 4K/16spp frame on one GPU (band ownership rank r of world N), for several
RTB200_CHUNK_RAYS values: python scripts/band_times.py [world=8] [chunk_rays ...]"""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import rtb200, scenes as S
    world = int(sys.argv[2])
    v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
    band = int(os.environ.get("BAND_ROWS", "8"))
    for rank in range(world):
        best = 1e9
        for it in range(6):
            img, st = sc.render(width=3840, height=2160, nb_ray=16, band_rows=band, rank=rank, world=world)
            best = min(best, st["kernel_ms"])
        img, st = sc.render(width=3840, height=2160, nb_ray=16, band_rows=band, rank=rank, world=world, profile=True)
        print("  band %d world %d rank %d chunk %s: overlapped %.3f ms | serial %.3f (primary %.3f shadow %.3f shade %.3f) launches %d" % (
            band, world, rank, os.environ.get("RTB200_CHUNK_RAYS", "default"), best, st["kernel_ms"], st["primary_ms"], st["shadow_ms"], st["shade_ms"], st["launches"]))
else:
    world = sys.argv[1] if len(sys.argv) > 1 else "8"
    for ch in (sys.argv[2:] or ["16777216", "8388608", "4194304", "2097152"]):
        env = dict(os.environ, RTB200_CHUNK_RAYS=ch)
        subprocess.run([sys.executable, __file__, "child", world], env=env)
