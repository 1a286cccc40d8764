"""Where a rank's end-to-end step goes (gpurun, one GPU): wall time per step of the band share `rank 0 of WORLD`, with and
without a new sample table per step (= with and without the light-grid rebuild), host frame in pinned memory.
usage: python scripts/e2e_breakdown.py [config 3|3b|4] [world ...]"""
import os, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import rtb200, scenes as S

cfg = sys.argv[1] if len(sys.argv) > 1 else "3"
worlds = [int(x) for x in sys.argv[2:]] or [1, 8]
if cfg == "4":
    v, c = S.instanced_bunnies(8); kw = dict(width=3840, height=2160, nb_ray=1)
else:
    v, c = S.big_bunny(); kw = dict(width=3840, height=2160, nb_ray=16, distance=576.0 if cfg == "3b" else 288.0)
H, W = kw["height"], kw["width"]
smp = torch.from_numpy(S.make_samples()).pin_memory().numpy()
smp2 = torch.from_numpy(S.make_samples(seed=0x5EED0002)).pin_memory().numpy()
out = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory().numpy()
sc = rtb200.Scene(v, c)
sc.set_samples(smp)
for world in worlds:
    rk = dict(kw, band_rows=4, rank=0, world=world, frame_rows=(world > 1)) if world > 1 else dict(kw)
    for _ in range(3):
        sc.render(out=out, **rk)

    def loop(n, new_table):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); km = []
        for i in range(n):
            if new_table:
                sc.set_samples_async(smp2 if i & 1 else smp)
            km.append(sc.render(out=out, **rk)[1]["kernel_ms"])
        return (time.perf_counter() - t0) / n * 1e3, float(np.mean(km))
    def pipelined(n):
        """begin / next table / end: the next frame's table and grids are prepared next to the frame in flight"""
        torch.cuda.synchronize()
        sc.set_samples_async(smp)
        t0 = time.perf_counter(); km = []
        for i in range(n):
            sc.render_begin(out, **rk)
            sc.set_samples_async(smp if i & 1 else smp2)
            km.append(sc.render_end()[1]["kernel_ms"])
        return (time.perf_counter() - t0) / n * 1e3, float(np.mean(km))
    loop(3, True)
    loop(4, False)                                   # (the third frame of a table rebuilds its grids finer, once)
    a, ka = loop(10, False)
    loop(3, True)                                    # (the first new table after that is built synchronously, once)
    b, kb = loop(10, True)
    pipelined(3)
    c, kc = pipelined(10)
    print("config %s world %d: pipelined (render_begin / set_samples_async / render_end) %.3f ms/step (events %.3f)" % (cfg, world, c, kc), flush=True)
    sc.render(out=out, **rk)
    st = sc.render(out=out, **rk)[1]
    print("config %s world %d: same table %.3f ms/step (events %.3f) | new table %.3f ms/step (events %.3f) | rebuild cost %.3f ms | redo %d crc %08x"
          % (cfg, world, a, ka, b, kb, b - a, st["grid_redo"], zlib.crc32(out.tobytes()) & 0xffffffff), flush=True)
