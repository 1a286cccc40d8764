"""Short GPU-only run for ncu: a few frames of one config through the C ABI (no oracle, no CPU work).
usage: python scripts/profile_frame.py [width height nb_ray frames [distance [instances]]]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S

a = sys.argv[1:]
w, h, nr, frames = (int(x) for x in (a[:4] + ["1920", "1080", "16", "3"][len(a[:4]):]))
dist = float(a[4]) if len(a) > 4 else 288.0
inst = int(a[5]) if len(a) > 5 else 1
v, c = S.instanced_bunnies(inst) if inst > 1 else S.big_bunny()
sc = rtb200.Scene(v, c)
sc.set_samples(S.make_samples())
for i in range(frames):
    img, st = sc.render(width=w, height=h, nb_ray=nr, distance=dist)
print(w, h, nr, dist, inst, st)
