"""Frame times of the BASELINE.json configs that are parity-test cases rather than bench lines (1, 2, 3 with distance 576, 4)
on one GPU: scene build, grid build, kernel time per frame (device-resident, best of 5), ray counts.
usage: python scripts/config_times.py"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtb200, scenes as S

smp = S.make_samples()
cases = [("1: bunny.obj 1920x1080 NB_RAY=1", S.small_bunny, dict(width=1920, height=1080, nb_ray=1)),
         ("2: big_bunny.obj 1920x1080 NB_RAY=1", S.big_bunny, dict(width=1920, height=1080, nb_ray=1)),
         ("3b: big_bunny.obj 3840x2160 NB_RAY=16 distance 576", S.big_bunny, dict(width=3840, height=2160, nb_ray=16, distance=576.0)),
         ("4: 64 big_bunny copies (317 953 primitives) 3840x2160 NB_RAY=1", S.instanced_bunnies, dict(width=3840, height=2160, nb_ray=1))]
for name, make, kw in cases:
    v, c = make()
    t0 = time.perf_counter()
    sc = rtb200.Scene(v, c); sc.set_samples(smp)
    build_s = time.perf_counter() - t0
    info = sc.info()
    best, st = 1e9, None
    for _ in range(5):
        img, s1 = sc.render(**kw)
        if s1["kernel_ms"] < best: best, st = s1["kernel_ms"], s1
    rays = st["primary_rays"] + st["shadow_rays"]
    print(json.dumps({"config": name, "primitives": info["n_tris"], "scene_create_s": round(build_s, 3), "bvh_build_ms": round(info["build_ms"], 3),
                      "grid_build_ms": round(st["grid_build_ms"], 3), "grid_path": st["grid_cells"] > 0, "kernel_ms_per_frame": round(best, 3),
                      "mrays_per_s_total": round(rays / best / 1e3, 1), "primary_hits": st["primary_hits"], "shadow_rays": st["shadow_rays"],
                      "shadow_occluded": st["shadow_occluded"], "queued_hit_points": st["queued_hits"], "launches": st["launches"]}), flush=True)
    del sc
