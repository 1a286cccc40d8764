"""Loader throughput: K x K flattened big_bunny copies as one OBJ (K = 64: 20.3 M triangles, ~0.8 GB of text) ->
serial reader vs the parallel reader -> pinned host array -> rtb_scene_create (one DMA + k_prepare_tris + BVH build).
usage: python scripts/loader_bench.py [K=64] [out json]"""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rtb200, scenes as S

K = int(sys.argv[1]) if len(sys.argv) > 1 else 64
tool = os.path.join(ROOT, "ray-tracer-rust_b200", "obj_instances")
path = "/tmp/bunnies_%d.obj" % K
t0 = time.perf_counter()
subprocess.check_call([tool, os.path.join(S.MODELS, "big_bunny.obj"), path, str(K)], stdout=subprocess.DEVNULL)
gen_s = time.perf_counter() - t0
size = os.path.getsize(path)
out = {"copies": K * K, "file_bytes": size, "generate_s": gen_s, "host_threads": os.cpu_count()}
if K <= 16:                                   # the serial robust reader, for the ratio (minutes at K = 64)
    t0 = time.perf_counter(); v_ser = rtb200.import_obj_ex(path, mtl=False)[0]; out["serial_ex_s"] = time.perf_counter() - t0
rtb200.import_obj_parallel(path, n_threads=0)          # warm the page cache
_w = rtb200.Scene(*S.random_soup(64, seed=1)); del _w   # CUDA context, module load: not part of any figure below
for pinned in (False, True):
    t0 = time.perf_counter()
    v, st = rtb200.import_obj_parallel(path, n_threads=0, pinned=pinned)
    wall = time.perf_counter() - t0
    key = "parallel_pinned" if pinned else "parallel_pageable"
    out[key] = {"wall_s": wall, "read_s": st["read_s"], "parse_s": st["parse_s"], "alloc_s": st["alloc_s"], "threads": st["threads"], "triangles": int(len(v)),
                "text_GBps": size / (st["read_s"] + st["parse_s"]) / 1e9, "triangles_per_s": len(v) / (st["read_s"] + st["parse_s"])}
    if K <= 16 and not pinned:
        out["parallel_equals_serial"] = bool((v.view(np.uint32) == v_ser.view(np.uint32)).all())
        del v_ser
    if not pinned:
        t0 = time.perf_counter(); sc0 = rtb200.Scene(v, None); out["scene_create_from_pageable_s"] = time.perf_counter() - t0
        del sc0, v
# scene creation from the pinned array: upload (36 B / triangle, one DMA) + k_prepare_tris + Morton sort + PLOC BVH; colours = NULL (white)
nbytes = v.nbytes
t0 = time.perf_counter()
sc = rtb200.Scene(v, None)
create_s = time.perf_counter() - t0
info = sc.info()
out["scene_create"] = {"wall_s": create_s, "upload_bytes": int(nbytes), "bvh_build_ms": info["build_ms"], "n_tris": info["n_tris"], "bvh_depth": info["max_depth"],
                       "upload_plus_build_GBps": nbytes / create_s / 1e9}
# the upload alone, for the PCIe figure
import torch
d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
h = torch.from_numpy(v.view(np.uint8).reshape(-1))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record(); d.copy_(h, non_blocking=True); e1.record(); torch.cuda.synchronize()
out["pinned_h2d_GBps"] = nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9
# a small frame of the loaded scene, against the numpy-instanced scene when it is small enough to build both
sc.set_samples(S.make_samples())
t0 = time.perf_counter()
img, st = sc.render(width=480, height=270, distance=72.0, nb_light_sample=4)
out["small_frame"] = {"wall_s": time.perf_counter() - t0, "kernel_ms": st["kernel_ms"], "primary_hits": st["primary_hits"], "path": "light grids" if st["grid_cells"] else "lists / walks"}
print(json.dumps(out))
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(json.dumps(out) + "\n")
os.remove(path)
