"""Small renders that cover every kernel of the path, for compute-sanitizer (memcheck / racecheck / initcheck):
  compute-sanitizer --tool memcheck python scripts/sanitize_run.py
Small chunks (two worker streams), the light-grid path and the list / walk path, the term wavefront, a coloured soup."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("RTB200_CHUNK_RAYS", "6000")
import numpy as np
import rtb200, scenes as S

smp = S.make_samples(20000)
v, c = S.big_bunny()
for mode in ("grid", "lists"):
    os.environ["RTB200_SHADOW"] = mode
    for terms in ("0", "1"):
        os.environ["RTB200_TERMS"] = terms
        sc = rtb200.Scene(v, c); sc.set_samples(smp)
        img, st = sc.render(width=160, height=90, distance=24.0, nb_ray=2, nb_light_sample=8)
        pix = S.stratified_pixels(160, 90, 9, seed=2)
        g = sc.render_pixels(pix, width=160, height=90, distance=24.0, nb_ray=2, nb_light_sample=8)
        print(mode, terms, st["launches"], st["queued_hits"], int(img.sum()), int(g["rgb8"].sum()))
        del sc
v, c = S.random_soup(300, seed=5)
os.environ["RTB200_SHADOW"] = "grid"; os.environ["RTB200_TERMS"] = "1"
sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples(7919, seed=5))
img, st = sc.render(width=150, height=83, nb_ray=2, nb_light_sample=7)
print("soup", st["launches"], int(img.sum()))
print("sanitize_run done")
