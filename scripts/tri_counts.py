"""tri_tests counter of the instrumented kernels, frame by frame (the light grids change form on the 2nd and 3rd frame of a table)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S
v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
kw = dict(width=3840, height=2160, nb_ray=16)
for k in range(5):
    _, st = sc.render(counters=True, **kw)
    _, s0 = sc.render(counters=True, **dict(kw, nb_light_sample=0))
    print(k, "full tri", st["tri_tests"], "box", st["box_tests"], "| primary only tri", s0["tri_tests"], "box", s0["box_tests"], "| cells", st["grid_cells"], "rebuilds", st["grid_rebuilds"], s0["grid_rebuilds"])
