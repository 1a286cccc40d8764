import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import rtb200, scenes as S
v, c = S.big_bunny(); sc = rtb200.Scene(v, c); sc.set_samples(S.make_samples())
for (w,h,nr) in ((1920,1080,1),(3840,2160,4)):
    img, st = sc.render(width=w, height=h, nb_ray=nr, counters=True)
    print(w,h,nr,st)
    c = sc.counters(); rays = st['queued_groups'] * 25
    print('  k2: occluded rays', c[8], 'steps/occluded ray', c[6] / max(c[8], 1), 'free rays ~', rays - c[8], 'steps/free ray', c[7] / max(rays - c[8], 1))
