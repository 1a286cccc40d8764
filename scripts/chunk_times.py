"""Per-chunk kernel times of one profiled 4K/16spp frame (warm): RTB200_TRACE_CHUNKS=1 python scripts/chunk_times.py"""
import os, sys
os.environ["RTB200_TRACE_CHUNKS"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S
v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
for i in range(3): sc.render(width=3840, height=2160, nb_ray=16)
sys.stderr.write("---- profiled frame\n"); sys.stderr.flush()
img, st = sc.render(width=3840, height=2160, nb_ray=16, profile=True)
print(st["primary_ms"], st["shadow_ms"], st["shade_ms"])
