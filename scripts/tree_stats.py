"""Dump the top of the device-built BVH and depth of selected primitives (development aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, rtb200, scenes as S
v, c = S.big_bunny()
sc = rtb200.Scene(v, c)
print(sc.info())
nodes, ch = sc.nodes()
depth = {}
stack = [(0, 0)]
leafdepth = np.zeros(len(v), int)
while stack:
    i, d = stack.pop()
    for k in range(2):
        c_ = ch[i, k]
        if c_ >= 0: stack.append((c_, d + 1))
        else: leafdepth[~c_] = d + 1
print("ground depth", leafdepth[-1], "mean leaf depth", leafdepth.mean(), "max", leafdepth.max())
def show(i, d, maxd):
    for k in range(2):
        b = nodes[i, 6 * k: 6 * k + 6]
        print("  " * d, "child", ch[i, k], "lo", b[:3].round(1), "hi", b[3:].round(1))
        if ch[i, k] >= 0 and d < maxd: show(ch[i, k], d + 1, maxd)
show(0, 0, 3)
