"""Independent pin of the oracle: complete render_pixel results (src/main.rs:180-240) derived step by step in numpy float32.

  python tests/golden/make_pixel_golden.py        # writes tests/golden/pixel_golden.json

This script shares NO code with oracle/rt_oracle.cpp or with the CUDA library: it restates the cited Rust lines directly in
numpy, one rounded float32 operation per numpy operation (numpy never fuses a*b+c), brute force over all 4 969 primitives of
big_bunny + ground for every ray (no tree: the reference's result is the minimum-distance leaf among those whose own box, own
Moller-Trumbore test and `x >= 1.0` rule pass -- bounding_volume_hierarchy.rs:50-142; ancestors' boxes are unions and cannot
reject what a leaf box accepts).  Pixels with an exact distance tie between two primitives are skipped (the tie order is the
only place the tree's shape matters), so no tree is needed here.

The fixture holds, per pixel: hit count, primitive and distance bits of ray 0, the linear RGB bits and the RGB8 bytes.
tests/test_oracle.py checks the C++ oracle against it on the CPU; tests/test_gpu_parity.py checks the CUDA path on the GPU.
Pixels are chosen by scanning candidates and keeping a fixed number per category (sky, lit ground, umbra, penumbra, lit bunny,
self-shadowed bunny, silhouette), at 1920x1080 / 1 ray per pixel and at 3840x2160 / 16 rays per pixel.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
F = np.float32
FLT_MAX = np.finfo(np.float32).max


def read_obj(path):                       # src/main.rs:114-149: `v` and `f` lines, single-space split, 1-based bare indices
    vs, tris = [], []
    for line in open(path).read().split("\n"):
        t = line.split(" ")
        if t[0] == "v":
            vs.append([F(t[1]), F(t[2]), F(t[3])])
        elif t[0] == "f":
            tris.append([vs[int(t[1]) - 1], vs[int(t[2]) - 1], vs[int(t[3]) - 1]])
    return np.array(tris, dtype=F)        # (n, 3, 3)


def samples_table(n=2000000, seed=0x5EED0001):      # include/rtb200.h rtb_fill_samples: SplitMix64, top 24 bits * 2^-24
    out = np.empty(2 * n, dtype=F)
    mask = (1 << 64) - 1
    state = seed
    # vectorised SplitMix64: state_j = seed + j * golden
    j = np.arange(1, 2 * n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + j * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    out[:] = (z >> np.uint64(40)).astype(F) * F(2.0 ** -24)
    del mask, state
    return out.reshape(n, 2)


def dot(a, b):                            # nalgebra 0.11 dot: left-to-right sum of products
    return (a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1]) + a[..., 2] * b[..., 2]


def cross(a, b):
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1],
                     a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], axis=-1)


def normalize(v):                         # Unit::new_normalize: v / sqrt(dot(v, v)) component-wise (ray.rs:12-17)
    n = np.sqrt(dot(v, v))
    return (v / (n[..., None] if np.ndim(n) else n)).astype(F), n


class Scene:
    def __init__(self):
        tri = read_obj(os.path.join(ROOT, "assets", "models", "big_bunny.obj"))
        ground = np.array([[[-10000.0, 0.0, -10000.0], [10000.0, 0.0, -10000.0], [0.0, 0.0, 10000.0]]], dtype=F)   # src/main.rs:102-111
        self.v = np.concatenate([tri, ground], axis=0)
        self.col = np.concatenate([np.ones((len(tri), 3), F), np.full((1, 3), 0.5, F)], axis=0)                    # :143, :108
        self.v0 = self.v[:, 0]
        self.e1 = (self.v[:, 1] - self.v[:, 0]).astype(F)         # triangle.rs:23-24
        self.e2 = (self.v[:, 2] - self.v[:, 0]).astype(F)
        self.normal, _ = normalize(cross(self.e1, self.e2))       # triangle.rs:29

        def mn(a, b):
            return np.where(a < b, a, b)                          # triangle.rs:37-39

        def mx(a, b):
            return np.where(a < b, b, a)                          # triangle.rs:41-43
        self.bmin = mn(mn(self.v[:, 0], self.v[:, 1]), self.v[:, 2])    # triangle.rs:45-56
        self.bmax = mx(mx(self.v[:, 0], self.v[:, 1]), self.v[:, 2])
        self.light = np.array([[-10.0, 300.0, -10.0], [10.0, 300.0, -10.0], [0.0, 300.0, 0.0]], dtype=F)           # src/main.rs:337-343

    def box_pass(self, o, d):             # bounding_box.rs:99-181 for every primitive's own box
        with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
            inside = np.all((o > self.bmin) & (o < self.bmax), axis=1)
            t = []
            for a in range(3):
                lo = ((self.bmin[:, a] - o[a]) / d[a]).astype(F)
                hi = ((self.bmax[:, a] - o[a]) / d[a]).astype(F)
                t.append((lo, hi) if d[a] >= F(0.0) else (hi, lo))
            tmin, tmax = t[0]
            tymin, tymax = t[1]
            rej1 = (tmin > tymax) | (tymin > tmax)
            tmin = np.where(tymin > tmin, tymin, tmin)
            tmax = np.where(tymax < tmax, tymax, tmax)
            tzmin, tzmax = t[2]
            rej2 = (tmin > tzmax) | (tzmin > tmax)
            tmin = np.where(tzmin > tmin, tzmin, tmin)
            tmax = np.where(tzmax < tmax, tzmax, tmax)
            ok = (~rej1) & (~rej2) & (tmin < FLT_MAX) & (tmax > F(0.0))
        return inside | ok

    def tri_t(self, o, d):                # triangle.rs:65-95 for every primitive: (accepted mask, distance)
        with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
            dd = np.broadcast_to(d, self.e2.shape)
            pvec = cross(dd, self.e2)
            det = dot(self.e1, pvec)
            par = (det < F(0.00001)) & (det > F(-0.00001))
            inv = (F(1.0) / det).astype(F)
            tvec = (o - self.v0).astype(F)
            u = dot(tvec, pvec) * inv
            bad_u = (u < F(0.0)) | (u > F(1.0))
            qvec = cross(tvec, self.e1)
            v = dot(dd, qvec) * inv
            bad_v = (v < F(0.0)) | ((u + v) > F(1.0))
            t = dot(self.e2, qvec) * inv
        return (~par) & (~bad_u) & (~bad_v), t.astype(F)

    def closest(self, o, d):
        """BoundingVolumeHierarchy::intersect: (prim, t) of the closest accepted leaf or (-1, 0); None on an exact tie."""
        ok, t = self.tri_t(o, d)
        ok = ok & self.box_pass(o, d) & ~(t < F(1.0))             # bounding_volume_hierarchy.rs:52, 64-67
        if not ok.any():
            return -1, F(0.0)
        tt = np.where(ok, t, np.inf)
        k = int(np.argmin(tt))
        if int((tt == tt[k]).sum()) != 1:
            return None
        return k, t[k]


def camera(eye, look_at, up):             # camera.rs:17-35
    eye, look_at, up = (np.array(x, dtype=F) for x in (eye, look_at, up))
    w, _ = normalize((eye - look_at).astype(F))
    o, _ = normalize(up)
    u, _ = normalize(cross(o, w))
    v, _ = normalize(cross(u, w))
    return u, v, w, eye


def light_point(tri, a, b):               # triangle.rs:113-127 (c3 = v * sqrt(u): sic)
    su, sv = np.sqrt(F(a)), np.sqrt(F(b))
    c1 = F(1.0) - su
    c2 = su * (F(1.0) - sv)
    c3 = F(b) * su
    return ((c1 * tri[0] + c2 * tri[1]) + c3 * tri[2]).astype(F)


def render_pixel(sc, cam, smp, px, py, width, height, nb_ray, nls, distance):
    """src/main.rs:151-240.  Returns dict or None when a tie was met."""
    u, v, w, eye = cam
    n_s = len(smp)
    avg = np.zeros(3, dtype=F)
    denom = F(nb_ray * nls)
    hit_count, prim0, t0, n_occ = 0, -1, F(0.0), 0
    for r in range(nb_ray):
        k = (px * width + py + r) % n_s                                           # :161 (px * width + py + i)
        A = (F(px) - F(width) / F(2.0)) + smp[k, 0]
        B = (F(py) - F(height) / F(2.0)) + smp[k, 1]
        direction = ((A * u + B * v).astype(F) - (F(distance) * w).astype(F)).astype(F)
        d, _ = normalize(direction)                                               # Ray::new
        hit = sc.closest(eye, d)
        if hit is None:
            return None
        prim, t = hit
        if prim < 0:
            continue
        hit_count += 1
        if r == 0:
            prim0, t0 = prim, t
        p_hit = (eye + (t * d).astype(F)).astype(F)                               # bounding_volume_hierarchy.rs:69
        col, nrm = sc.col[prim], sc.normal[prim]
        for i in range(nls):
            a, b = smp[(r * nb_ray + i) % n_s]                                    # :194
            p = light_point(sc.light, a, b)
            sdir = (p - p_hit).astype(F)
            sd, _ = normalize(sdir)                                               # :201
            diff = (p - p_hit).astype(F)
            L = np.sqrt(dot(diff, diff))                                          # :202 distance(p, origin)
            h2 = sc.closest(p_hit, sd)                                            # :204
            if h2 is None:
                return None
            lnd = np.abs(dot(nrm, sd))                                            # :206-207
            lit = True
            if h2[0] >= 0:
                x_hit = (p_hit + (h2[1] * sd).astype(F)).astype(F)
                dd = (p_hit - x_hit).astype(F)
                lit = bool(np.sqrt(dot(dd, dd)) > L)                              # :220
            if lit:
                avg = (avg + ((col * lnd).astype(F) / denom).astype(F)).astype(F)  # :209-216
            else:
                avg = (avg + ((np.zeros(3, F) * F(1.0)) / denom).astype(F)).astype(F)
                n_occ += 1
    g = np.power(avg, F(1.0) / F(2.2)).astype(F) * F(255.0)                       # color.rs:10-13, 26-33
    rgb8 = [int(x) if x == x else 0 for x in np.clip(np.trunc(g), 0, 255)]
    return dict(px=int(px), py=int(py), hit_count=hit_count, prim0=int(prim0), t0_bits=int(np.array(t0, F).view(np.uint32)),
                rgb_lin_bits=[int(x) for x in avg.view(np.uint32)], rgb8=rgb8, occluded=n_occ)


def classify(res, n_bunny, nls, nb_ray):
    if res["hit_count"] == 0:
        return "sky"
    total = res["hit_count"] * nls
    if res["hit_count"] < nb_ray:
        return "silhouette"
    on_bunny = 0 <= res["prim0"] < n_bunny
    if res["occluded"] == 0:
        return "bunny_lit" if on_bunny else "ground_lit"
    if res["occluded"] == total:
        return "bunny_dark" if on_bunny else "umbra"
    return "bunny_partial" if on_bunny else "penumbra"


def main():
    sc = Scene()
    smp = samples_table()
    cam = camera((0.0, 100.0, 200.0), (0.0, 0.0, -100000.0), (0.0, 1.0, 0.0))     # src/main.rs:353-355
    n_bunny = len(sc.v) - 1
    out = {"scene": "big_bunny.obj + ground (src/main.rs:102-111), light src/main.rs:337-343, camera src/main.rs:353-356",
           "samples": "SplitMix64 seed 0x5EED0001, 2000000 pairs", "configs": []}
    rng = np.random.default_rng(20261105)
    plans = [
        dict(width=1920, height=1080, nb_ray=1, nls=100, distance=288.0,
             want={"sky": 2, "ground_lit": 4, "umbra": 3, "penumbra": 6, "bunny_lit": 3, "bunny_partial": 5, "bunny_dark": 1},
             window=(780, 1040, 430, 800), extra=[(5, 5), (1919, 1079), (960, 1079), (3, 800)]),
        dict(width=3840, height=2160, nb_ray=16, nls=100, distance=288.0,
             want={"ground_lit": 1, "penumbra": 2, "bunny_partial": 1, "silhouette": 2, "bunny_lit": 1},
             window=(1600, 2060, 880, 1560), extra=[(3000, 2000), (1900, 1082), (2500, 1082), (1840, 1000), (2100, 1082), (2100, 1083), (1620, 1082), (1620, 1083), (2040, 1083)] + [(1800, y) for y in range(880, 1000, 7)]),
    ]
    for plan in plans:
        want = dict(plan["want"])
        got = []
        x0, x1, y0, y1 = plan["window"]
        cands = list(plan["extra"]) + [(int(rng.integers(x0, x1)), int(rng.integers(y0, y1))) for _ in range(400)]
        for (px, py) in cands:
            if not any(want.values()):
                break
            res = render_pixel(sc, cam, smp, px, py, plan["width"], plan["height"], plan["nb_ray"], plan["nls"], plan["distance"])
            if res is None:
                continue
            c = classify(res, n_bunny, plan["nls"], plan["nb_ray"])
            if want.get(c, 0) > 0 or (px, py) in plan["extra"]:
                if want.get(c, 0) > 0:
                    want[c] -= 1
                res["category"] = c
                got.append(res)
                print(plan["width"], plan["nb_ray"], c, res["px"], res["py"], res["hit_count"], res["occluded"], res["rgb8"], file=sys.stderr, flush=True)
        out["configs"].append({k: plan[k] for k in ("width", "height", "nb_ray", "nls", "distance")} | {"pixels": got, "unfilled": {k: v for k, v in want.items() if v}})
    json.dump(out, open(os.path.join(HERE, "pixel_golden.json"), "w"), indent=1)
    print("pixels:", [len(c["pixels"]) for c in out["configs"]])


if __name__ == "__main__":
    main()
