import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import rtb200, scenes as S
v, c = S.instanced_bunnies()
sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
for i in range(2):
    img, st = sc.render(width=3840, height=2160, nb_ray=1, profile=True)
print({k: st[k] for k in ("kernel_ms", "primary_ms", "shadow_ms", "shade_ms", "queued_hits", "queued_groups", "queued_walks", "grid_cells", "launches")})
