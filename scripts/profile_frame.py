"""Short GPU-only run for ncu: a few frames of one config through the C ABI (no oracle, no CPU work).
usage: python scripts/profile_frame.py [width height nb_ray frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S

w, h, nr, frames = (int(x) for x in (sys.argv[1:5] + ["1920", "1080", "16", "3"][len(sys.argv) - 1:]))
v, c = S.big_bunny()
sc = rtb200.Scene(v, c)
sc.set_samples(S.make_samples())
for i in range(frames):
    img, st = sc.render(width=w, height=h, nb_ray=nr)
print(w, h, nr, st)
