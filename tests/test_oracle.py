"""CPU tests of the oracle (oracle/rt_oracle.cpp) against the hand-derived known answers and against itself
(reference-tree traversal == brute force == any-hit form).  "parity unpinned": the reference has no tests."""
import json
import os

import numpy as np
import pytest

import scenes as S
from oracle import oracle as O

KA = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "known_answers.json")))


def test_camera_basis_known_answer(built):
    cam = KA["camera_default"]
    u, v, w = O.camera_basis(cam["eye"], cam["look_at"], cam["up"])
    for name, got in (("u", u), ("v", v), ("w", w)):
        assert np.allclose(got.astype(np.float64), cam[name], rtol=cam["rtol"], atol=1e-12), name
    # v = u x w points down (image y grows downwards), basis is orthonormal
    assert v[1] < 0
    assert abs(float(np.dot(u, v))) < 1e-6 and abs(float(np.dot(v, w))) < 1e-6


def test_light_sample_known_answers(built):
    for e in KA["light_sample"]:
        p = O.light_sample(S.LIGHT, *e["uv"])
        assert [float(x) for x in p] == e["p"], e
    # c3 uses v, not sqrt(v): the weights do not sum to one, samples leave the light's plane
    assert O.light_sample(S.LIGHT, 0.25, 0.25)[1] != 300.0


def test_gamma_known_answers_and_monotone(built):
    for c, b in KA["gamma_u8"]:
        assert O.gamma_u8(c) == b
    assert O.gamma_u8(float("nan")) == 0 and O.gamma_u8(-1.0) == 0 and O.gamma_u8(7.0) == 255
    xs = np.sort(np.random.default_rng(0).uniform(0, 1.05, 20000).astype(np.float32))
    bs = np.array([O.gamma_u8(x) for x in xs])
    assert (np.diff(bs) >= 0).all()


def test_box_known_answers(built):
    for e in KA["box"]:
        assert O.box_intersect(e["min"], e["max"], e["o"], e["d"]) == e["hit"], e["why"]


def test_tri_known_answers(built):
    for e in KA["tri"]:
        t = O.tri_intersect(e["v"], e["o"], e["d"])
        if e["t"] is None:
            assert t is None, e
        else:
            assert t == pytest.approx(e["t"], abs=1e-6), e


def test_reference_tree_shape(oracle_big_bunny):
    # one leaf per primitive, binary: 2n-1 nodes (bounding_volume_hierarchy.rs:173-226)
    assert oracle_big_bunny.n == 4969
    assert oracle_big_bunny.num_nodes() == 2 * 4969 - 1
    r = oracle_big_bunny.dfs_rank()
    assert sorted(r.tolist()) == list(range(4969))


def test_ground_pixel_known_answer(oracle_big_bunny):
    g = KA["ground_pixel"]
    zero = np.zeros((1000, 2), np.float32)
    r = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, zero, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT,
                                pixels=[[g["px"], g["py"]]])
    assert r["hit_count"][0] == 1 and r["prim0"][0] == oracle_big_bunny.n - 1
    assert r["t0"][0] == pytest.approx(g["t"], abs=g["t_atol"])
    assert np.allclose(oracle_big_bunny.normal(r["prim0"][0]), g["normal"])
    # all 100 samples are the light's v0 (u=v=0): unoccluded grey ground, lnd = |n.l| < 1
    assert 0.3 < r["rgb_lin"][0, 0] < 0.5 and r["rgb_lin"][0, 0] == r["rgb_lin"][0, 1] == r["rgb_lin"][0, 2]


def _same(a, b):
    for k in ("hit_count", "prim0", "rgb8"):
        assert (a[k] == b[k]).all(), k
    assert (a["t0"].view(np.uint32) == b["t0"].view(np.uint32)).all()
    assert (a["rgb_lin"].view(np.uint32) == b["rgb_lin"].view(np.uint32)).all()


def test_tree_equals_brute_force_big_bunny(oracle_big_bunny, big_bunny_inputs):
    """the reference tree's result is topology-free -- brute force over primitives with the leaf-box
    test gives the same bits; and the any-hit shadow form gives the same colours."""
    _, _, smp = big_bunny_inputs
    pix = S.stratified_pixels(S.WIDTH, S.HEIGHT, 60, seed=5)
    args = (S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT)
    a = oracle_big_bunny.render(*args, pixels=pix, nb_light_sample=20, mode=O.MODE_REFERENCE_TREE)
    b = oracle_big_bunny.render(*args, pixels=pix, nb_light_sample=20, mode=O.MODE_BRUTE)
    c = oracle_big_bunny.render(*args, pixels=pix, nb_light_sample=20, mode=O.MODE_BRUTE_ANYHIT)
    _same(a, b)
    _same(a, c)
    st = a["stats"]
    assert st["primary_hits"] > 0 and st["shadow_occluded"] > 0
    assert st["tri_tests"] == b["stats"]["tri_tests"]      # a leaf is reached iff its own box passes
    assert st["nan_asserts"] == 0


def test_tree_equals_brute_force_soup(built):
    v, c = S.random_soup(300, seed=11)
    sc = O.OracleScene(v, c, True)
    smp = S.make_samples(5000, seed=7)
    args = (160, 90, smp, S.EYE, S.LOOK_AT, S.UP, 60.0, S.LIGHT)
    a = sc.render(*args, nb_ray=2, nb_light_sample=8, mode=O.MODE_REFERENCE_TREE)
    b = sc.render(*args, nb_ray=2, nb_light_sample=8, mode=O.MODE_BRUTE)
    c2 = sc.render(*args, nb_ray=2, nb_light_sample=8, mode=O.MODE_BRUTE_ANYHIT)
    _same(a, b)
    _same(a, c2)
    assert a["stats"]["primary_hits"] > 1000


def test_image_statistics_big_bunny(oracle_big_bunny, big_bunny_inputs):
    """about half of the primary rays hit (ground 48 %, bunny 2 %), ~7 % of shadow rays occluded."""
    _, _, smp = big_bunny_inputs
    pix = S.stratified_pixels(S.WIDTH, S.HEIGHT, 40, seed=2)
    r = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix,
                                nb_light_sample=10)
    st = r["stats"]
    assert 0.45 < st["primary_hits"] / st["primary_rays"] < 0.55
    assert 0.03 < st["shadow_occluded"] / st["shadow_rays"] < 0.12
    bunny = (r["prim0"] >= 0) & (r["prim0"] < oracle_big_bunny.n - 1)
    assert 0.005 < bunny.mean() < 0.04


def test_sample_index_is_px_times_width_plus_py(built):
    """create_rays indexes the table with px*width + py (src/main.rs:162), not py*width + px."""
    v, c = S.GROUND, S.GROUND_COLOR
    sc = O.OracleScene(v, c, True)
    smp = np.zeros((100000, 2), np.float32)
    W, H = 64, 48
    px, py = 10, 40
    smp[(px * W + py) % len(smp)] = (0.5, 0.5)
    base = sc.render(W, H, np.zeros_like(smp), S.EYE, S.LOOK_AT, S.UP, 20.0, S.LIGHT, pixels=[[px, py]], nb_light_sample=1)
    moved = sc.render(W, H, smp, S.EYE, S.LOOK_AT, S.UP, 20.0, S.LIGHT, pixels=[[px, py]], nb_light_sample=1)
    assert base["hit_count"][0] == 1 and moved["hit_count"][0] == 1
    assert base["t0"][0] != moved["t0"][0]


def _golden():
    return json.load(open(os.path.join(os.path.dirname(__file__), "golden", "pixel_golden.json")))


def check_against_pixel_golden(render_pixels):
    """Shared by the CPU (oracle) and GPU (CUDA path) tests: complete render_pixel results derived in numpy by
    tests/golden/make_pixel_golden.py, which shares no code with the oracle or the library."""
    g = _golden()
    n_total = 0
    for cfg in g["configs"]:
        px = np.array([[p["px"], p["py"]] for p in cfg["pixels"]], np.uint32)
        got = render_pixels(px, cfg)
        for i, p in enumerate(cfg["pixels"]):
            what = "%dx%d nb_ray %d pixel (%d,%d) %s" % (cfg["width"], cfg["height"], cfg["nb_ray"], p["px"], p["py"], p["category"])
            assert int(got["hit_count"][i]) == p["hit_count"], what
            assert int(got["prim0"][i]) == p["prim0"], what
            assert int(got["t0"][i:i + 1].view(np.uint32)[0]) == p["t0_bits"], what
            assert [int(x) for x in got["rgb_lin"][i].view(np.uint32)] == p["rgb_lin_bits"], what
            assert [int(x) for x in got["rgb8"][i]] == p["rgb8"], what
        n_total += len(cfg["pixels"])
    cats = {p["category"] for cfg in g["configs"] for p in cfg["pixels"]}
    assert n_total >= 20 and {"ground_lit", "umbra", "penumbra", "bunny_lit", "bunny_partial", "silhouette"} <= cats
    return n_total


def test_oracle_against_independent_numpy_pixels(oracle_big_bunny, big_bunny_inputs):
    """The pin of the oracle: whole pixels (primary hit + 100 shadow rays + ordered accumulate + gamma) bit for bit."""
    _, _, smp = big_bunny_inputs

    def render(px, cfg):
        return oracle_big_bunny.render(cfg["width"], cfg["height"], smp, S.EYE, S.LOOK_AT, S.UP, cfg["distance"], S.LIGHT,
                                       nb_ray=cfg["nb_ray"], nb_light_sample=cfg["nls"], pixels=px)
    check_against_pixel_golden(render)
    # and in the oracle's other traversal mode
    def render_brute(px, cfg):
        return oracle_big_bunny.render(cfg["width"], cfg["height"], smp, S.EYE, S.LOOK_AT, S.UP, cfg["distance"], S.LIGHT,
                                       nb_ray=cfg["nb_ray"], nb_light_sample=cfg["nls"], pixels=px, mode=O.MODE_BRUTE)
    check_against_pixel_golden(render_brute)
