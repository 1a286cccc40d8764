"""Round statistics of k_shadow_grid (build with -DRTB_DIAG_GRID=1 -DRTB_NUM_CTRS=32; RTB200_LIB=variants/librtb200_diag2.so)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S
v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
img, st = sc.render(width=3840, height=2160, nb_ray=16, counters=True)
c = sc.counters()
tk = max(int(c[16]), 1); rounds = max(int(c[18]), 1)
print("tickets", c[16], "queued rays per ticket", c[17] / tk, "test rounds per ticket", c[18] / tk, "avg queue length at round start", c[19] / rounds)
print("rounds by queue length: >=32 %d, 17-31 %d, 9-16 %d, 5-8 %d, 3-4 %d, 1-2 %d" % tuple(int(x) for x in c[20:26]))
print("lane slots with a candidate", c[26], "hits", c[27], "(%.3f)" % (c[27] / max(int(c[26]), 1)), "skipped / early-rejected without a t", c[28])
print("confirm batches per ticket", c[29] / tk, "rays per batch", c[30] / max(int(c[29]), 1), "tri_tests", st["tri_tests"], "occluded", st["shadow_occluded"])
