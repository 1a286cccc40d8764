"""Shared scene/input builders for the tests and bench.py (harness code, numpy only).

The literals restate the reference's main(): ground (src/main.rs:102-111), light
(src/main.rs:337-347), camera (src/main.rs:353-356), resolution (src/main.rs:350-351),
NB_RAY / NB_LIGHT_SAMPLE / NB_RAND_SAMPLE (src/main.rs:38-40).
"""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODELS = os.path.join(ROOT, "assets", "models")

GROUND = np.array([[-10000.0, 0.0, -10000.0, 10000.0, 0.0, -10000.0, 0.0, 0.0, 10000.0]], dtype=np.float32)
GROUND_COLOR = np.array([[0.5, 0.5, 0.5]], dtype=np.float32)
LIGHT = np.array([-10.0, 300.0, -10.0, 10.0, 300.0, -10.0, 0.0, 300.0, 0.0], dtype=np.float32)
EYE = (0.0, 100.0, 200.0)
LOOK_AT = (0.0, 0.0, -100000.0)
UP = (0.0, 1.0, 0.0)
DISTANCE = 288.0
WIDTH, HEIGHT = 1920, 1080
NB_RAY, NB_LIGHT_SAMPLE, NB_RAND_SAMPLE = 1, 100, 2000000
SAMPLE_SEED = 0x5EED0001


def parse_obj(path):
    """src/main.rs:114-149: only `v x y z` and `f i j k` lines (split on single spaces), 1-based bare
    indices, every triangle white.  Returns (n,9) float32."""
    verts, tris = [], []
    with open(path, "r") as f:
        for line in f.read().split("\n"):
            tok = line.split(" ")
            if tok[0] == "v":
                verts.append((np.float32(tok[1]), np.float32(tok[2]), np.float32(tok[3])))
            elif tok[0] == "f":
                i, j, k = int(tok[1]), int(tok[2]), int(tok[3])
                tris.append(verts[i - 1] + verts[j - 1] + verts[k - 1])
    return np.array(tris, dtype=np.float32).reshape(-1, 9)


def scene_from_objs(paths):
    """main(): OBJ triangles in argv order (white), then the ground triangle (src/main.rs:324-335)."""
    parts = [parse_obj(p) for p in paths]
    parts = [p for p in parts if len(p)]
    verts = np.concatenate(parts + [GROUND], axis=0)
    colors = np.concatenate([np.ones((len(verts) - 1, 3), np.float32), GROUND_COLOR], axis=0)
    return np.ascontiguousarray(verts), np.ascontiguousarray(colors)


def big_bunny():
    return scene_from_objs([os.path.join(MODELS, "big_bunny.obj")])


def small_bunny():
    return scene_from_objs([os.path.join(MODELS, "bunny.obj")])


def instanced_bunnies(k=8):
    """BASELINE config 4: k*k flattened copies, copy (i,j) translated by ((i-(k-1)/2)*200, 0, -j*200) in f32."""
    base = parse_obj(os.path.join(MODELS, "big_bunny.obj")).reshape(-1, 3, 3)
    parts = []
    for i in range(k):
        for j in range(k):
            off = np.array([(i - (k - 1) / 2.0) * 200.0, 0.0, -j * 200.0], dtype=np.float32)
            parts.append((base + off).astype(np.float32).reshape(-1, 9))
    verts = np.concatenate(parts + [GROUND], axis=0)
    colors = np.concatenate([np.ones((len(verts) - 1, 3), np.float32), GROUND_COLOR], axis=0)
    return np.ascontiguousarray(verts), np.ascontiguousarray(colors)


def random_soup(n, seed, extent=100.0, size=12.0, with_ground=True):
    """Seeded triangle soup in front of the default camera, random colours (parity stress scene)."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, size=(n, 1, 3)).astype(np.float32)
    c[:, 0, 1] = np.abs(c[:, 0, 1]) + 5.0
    c[:, 0, 2] -= 150.0
    v = (c + rng.uniform(-size, size, size=(n, 3, 3)).astype(np.float32)).astype(np.float32).reshape(n, 9)
    col = rng.uniform(0.05, 1.0, size=(n, 3)).astype(np.float32)
    if with_ground:
        v = np.concatenate([v, GROUND], axis=0)
        col = np.concatenate([col, GROUND_COLOR], axis=0)
    return np.ascontiguousarray(v), np.ascontiguousarray(col)


def make_samples(n=NB_RAND_SAMPLE, seed=SAMPLE_SEED):
    """Deterministic stand-in for the reference's thread_rng table (src/main.rs:260-265): SplitMix64 stream,
    value j = top 24 bits of output j times 2^-24 (uniform in [0,1), exactly representable); pair k = (2k, 2k+1)."""
    with np.errstate(over="ignore"):
        idx = np.arange(1, 2 * n + 1, dtype=np.uint64)
        z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return ((z >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)).reshape(n, 2)


def orbit_pose(k, n=120):
    """BASELINE config 5 pose k (pose 0 == reference camera)."""
    th = 2.0 * np.pi * k / n
    s, c = np.sin(th), np.cos(th)
    return (200.0 * s, 100.0, 200.0 * c), (-100000.0 * s, 0.0, -100000.0 * c), UP


def stratified_pixels(width, height, step, seed=1):
    """One random pixel per step x step cell -> (m,2) uint32 (px,py)."""
    rng = np.random.default_rng(seed)
    xs = np.arange(0, width, step)
    ys = np.arange(0, height, step)
    gx, gy = np.meshgrid(xs, ys, indexing="xy")
    px = np.minimum(gx + rng.integers(0, step, gx.shape), width - 1)
    py = np.minimum(gy + rng.integers(0, step, gy.shape), height - 1)
    return np.stack([px.ravel(), py.ravel()], axis=1).astype(np.uint32)
