"""Diagnostics of k_shadow_grid's round structure (COUNT build of the kernels): python scripts/diag_grid.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S
v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
for (w, h, nr) in ((1920, 1080, 1), (3840, 2160, 16)):
    img, st = sc.render(width=w, height=h, nb_ray=nr, counters=True)
    c = sc.counters()
    tk = max(c[12], 1); rounds = max(c[6], 1)
    print(w, h, nr, 'queued', st['queued_hits'], 'tri_tests', st['tri_tests'])
    print('  tickets', c[12], 'rays with a non-empty cell per ticket', c[14] / tk, 'entries per such ray', c[13] / max(c[14], 1))
    print('  rounds per ticket', c[6] / tk, 'active rays per round', c[7] / rounds, 'hits per round', c[15] / rounds)
