"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on the same seeded
inputs.  Bars: hit masks, hit primitive, hit distance bits, linear colour bits and RGB8 bytes all EXACT (the
north_star allows +-1 LSB on 0.1 % of channels and grazing-ray mask differences; this design needs neither)."""
import os

import numpy as np
import pytest

import scenes as S
from oracle import oracle as O

pytestmark = pytest.mark.gpu

CAM = dict(eye=S.EYE, look_at=S.LOOK_AT, up=S.UP)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def rtb(built):
    import rtb200
    return rtb200


@pytest.fixture(scope="module")
def gpu_big_bunny(rtb, big_bunny_inputs):
    v, c, smp = big_bunny_inputs
    sc = rtb.Scene(v, c)
    sc.set_samples(smp)
    return sc


def assert_same(g, o, what=""):
    assert (g["hit_count"] == o["hit_count"]).all(), what + " hit mask"
    assert (g["prim0"] == o["prim0"]).all(), what + " hit primitive"
    assert (g["t0"].view(np.uint32) == o["t0"].view(np.uint32)).all(), what + " hit distance bits"
    assert (g["rgb_lin"].view(np.uint32) == o["rgb_lin"].view(np.uint32)).all(), what + " linear colour bits"
    assert (g["rgb8"] == o["rgb8"]).all(), what + " rgb8"
    for k in ("primary_rays", "primary_hits", "shadow_rays", "shadow_occluded"):
        assert g["stats"][k] == o["stats"][k], what + " " + k


def test_scene_info(gpu_big_bunny):
    info = gpu_big_bunny.info()
    assert info["n_tris"] == 4969 and info["n_nodes"] == 4968
    assert 10 <= info["max_depth"] < 64


def test_big_bunny_1080p_subset_exact(gpu_big_bunny, oracle_big_bunny, big_bunny_inputs):
    """BASELINE config 2 on a stratified 1/400 pixel subset, reference-tree oracle."""
    _, _, smp = big_bunny_inputs
    pix = S.stratified_pixels(S.WIDTH, S.HEIGHT, 20, seed=9)
    g = gpu_big_bunny.render_pixels(pix, counters=True)
    o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
    assert_same(g, o, "big_bunny 1080p")
    assert g["stats"]["tri_tests"] > 0          # (box tests: none on the camera-grid + light-grid path)
    assert g["stats"]["box_tests"] < o["stats"]["box_tests"] / 20     # the point of building our own tree


def test_bunny_region_dense_exact(gpu_big_bunny, oracle_big_bunny, big_bunny_inputs):
    """Every pixel of a window over the bunny's head and its shadow edge (silhouettes, grazing rays)."""
    _, _, smp = big_bunny_inputs
    xs, ys = np.meshgrid(np.arange(800, 960, 2), np.arange(440, 520, 2))
    pix = np.stack([xs.ravel(), ys.ravel()], 1).astype(np.uint32)
    g = gpu_big_bunny.render_pixels(pix)
    o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix,
                                nb_light_sample=100)
    assert_same(g, o, "bunny window")
    assert ((g["prim0"] >= 0) & (g["prim0"] < 4968)).mean() > 0.3


def test_frame_kernel_equals_list_kernel_and_oracle(gpu_big_bunny, oracle_big_bunny, big_bunny_inputs):
    _, _, smp = big_bunny_inputs
    frame, st = gpu_big_bunny.render()
    assert frame.shape == (S.HEIGHT, S.WIDTH, 3)
    pix = S.stratified_pixels(S.WIDTH, S.HEIGHT, 24, seed=4)
    o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
    assert (frame[pix[:, 1], pix[:, 0]] == o["rgb8"]).all()
    assert st["primary_rays"] == S.WIDTH * S.HEIGHT
    assert 0.45 < st["primary_hits"] / st["primary_rays"] < 0.55
    assert st["shadow_rays"] == st["primary_hits"] * 100
    # sky is black (image zero-initialised, src/main.rs:247), ground is lit
    assert frame[:400].max() == 0 and frame[900:, :, :].min() > 100
    # determinism: the reference is not (thread_rng); this path must be
    frame2, _ = gpu_big_bunny.render()
    assert (frame == frame2).all()


def test_4k_16spp_subset_exact(gpu_big_bunny, oracle_big_bunny, big_bunny_inputs):
    """BASELINE config 3 (the bench workload) on a pixel subset: 16 rays/pixel, light-sample indices up to 339."""
    _, _, smp = big_bunny_inputs
    rng = np.random.default_rng(1)
    pix = np.stack([rng.integers(0, 3840, 150), rng.integers(900, 2160, 150)], 1).astype(np.uint32)
    pix = np.concatenate([pix, [[1800 + i, 1000 + j] for i in range(0, 200, 20) for j in range(0, 200, 20)]]).astype(np.uint32)
    g = gpu_big_bunny.render_pixels(pix, width=3840, height=2160, nb_ray=16)
    o = oracle_big_bunny.render(3840, 2160, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix, nb_ray=16)
    assert_same(g, o, "4K 16spp")
    assert g["hit_count"].max() == 16


def test_small_bunny_config1_exact(rtb, built):
    """BASELINE config 1: bunny.obj is sub-pixel at the default camera and many of its triangles fail the absolute
    |det| < 1e-5 test."""
    v, c = S.small_bunny()
    smp = S.make_samples()
    sc = rtb.Scene(v, c)
    sc.set_samples(smp)
    osc = O.OracleScene(v, c, True)
    pix = np.concatenate([S.stratified_pixels(S.WIDTH, S.HEIGHT, 40, seed=3),
                          [[x, y] for x in range(955, 966) for y in range(880, 900)]]).astype(np.uint32)
    g = sc.render_pixels(pix)
    o = osc.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
    assert_same(g, o, "bunny.obj")


@pytest.mark.parametrize("seed,n,spp,nls", [(1, 200, 1, 16), (2, 1000, 3, 7), (3, 37, 2, 100), (4, 2, 1, 5)])
def test_random_soup_full_frames_exact(rtb, built, seed, n, spp, nls):
    """Coloured triangle soups (interpenetrating, sliver and back-facing triangles), full small frames whose
    size is not a multiple of the 32x8 block tile."""
    v, c = S.random_soup(n, seed=seed)
    smp = S.make_samples(7919, seed=seed)              # short table: exercises the modulo wrap
    W, H = 150, 83
    sc = rtb.Scene(v, c)
    sc.set_samples(smp)
    osc = O.OracleScene(v, c, True)
    o = osc.render(W, H, smp, S.EYE, S.LOOK_AT, S.UP, 50.0, S.LIGHT, nb_ray=spp, nb_light_sample=nls)
    frame, st = sc.render(width=W, height=H, distance=50.0, nb_ray=spp, nb_light_sample=nls)
    assert (frame == o["rgb8"]).all()
    ys, xs = np.mgrid[0:H, 0:W]
    pix = np.stack([xs.ravel(), ys.ravel()], 1).astype(np.uint32)
    g = sc.render_pixels(pix, width=W, height=H, distance=50.0, nb_ray=spp, nb_light_sample=nls)
    flat = {k: (o[k].reshape(-1, 3) if o[k].ndim == 3 else o[k].ravel()) for k in ("rgb_lin", "rgb8", "hit_count", "prim0", "t0")}
    flat["stats"] = o["stats"]
    assert_same(g, flat, "soup %d" % seed)


def test_zero_jitter_axis_aligned_rays(gpu_big_bunny, rtb, oracle_big_bunny, big_bunny_inputs):
    """All-zero sample table: column px = width/2 has d.x == 0 exactly (inf/NaN slab arithmetic in the reference);
    those rays take the cull-nothing path here and must still agree."""
    v, c, _ = big_bunny_inputs
    zero = np.zeros((1000, 2), np.float32)
    sc = rtb.Scene(v, c)
    sc.set_samples(zero)
    pix = np.array([[960, y] for y in range(0, 1080, 9)] + [[959, 700], [961, 700], [0, 0], [1919, 1079]], np.uint32)
    g = sc.render_pixels(pix, nb_light_sample=3)
    o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, zero, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix, nb_light_sample=3)
    assert_same(g, o, "zero jitter")
    assert o["stats"]["nan_asserts"] == 0


def test_orbit_poses_exact(gpu_big_bunny, oracle_big_bunny, big_bunny_inputs):
    """BASELINE config 5 poses (camera-only change, scene stays resident)."""
    _, _, smp = big_bunny_inputs
    pix = S.stratified_pixels(S.WIDTH, S.HEIGHT, 60, seed=8)
    for k in (0, 17, 60, 95):
        eye, look, up = S.orbit_pose(k)
        g = gpu_big_bunny.render_pixels(pix, eye=eye, look_at=look, up=up, nb_light_sample=20)
        o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, eye, look, up, S.DISTANCE, S.LIGHT, pixels=pix, nb_light_sample=20)
        assert_same(g, o, "orbit pose %d" % k)


def test_instanced_bunnies_subset(rtb, built):
    """BASELINE config 4 scene (64 copies, 317 953 primitives); oracle in brute-force mode (no reference tree:
    O(n^2) build), tie rule = primitive index on both sides."""
    v, c = S.instanced_bunnies(8)
    assert len(v) == 317953
    smp = S.make_samples()
    sc = rtb.Scene(v, c)
    sc.set_samples(smp)
    assert (sc.tie_ranks() == np.arange(len(v))).all()
    rng = np.random.default_rng(5)
    pix = np.stack([rng.integers(0, 3840, 120), rng.integers(700, 2160, 120)], 1).astype(np.uint32)
    g = sc.render_pixels(pix, width=3840, height=2160, nb_light_sample=6)
    o = O.OracleScene(v, c, False).render(3840, 2160, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix,
                                          nb_light_sample=6, mode=O.MODE_BRUTE)
    assert_same(g, o, "64 bunnies")
    assert (g["prim0"] >= 0).sum() > 60


def test_tie_ranks_replay_reference_tree(gpu_big_bunny, oracle_big_bunny, rtb, built):
    assert (gpu_big_bunny.tie_ranks() == oracle_big_bunny.dfs_rank()).all()
    for seed, n in ((1, 2), (2, 3), (3, 64), (4, 1025), (5, 3000)):
        v, c = S.random_soup(n, seed=seed, with_ground=(seed % 2 == 0))
        assert (rtb.Scene(v, c).tie_ranks() == O.OracleScene(v, c, True).dfs_rank()).all(), n


def test_exact_distance_ties_follow_reference_leaf_order(rtb, built):
    """Coincident triangles with different colours: equal t bit for bit, the later leaf of the reference's tree wins
    (bounding_volume_hierarchy.rs:123-130)."""
    base = np.array([-40, 20, -60, 40, 20, -60, 0, 90, -60], np.float32)
    v = np.stack([base, base, base, base + np.float32(500.0)] + [S.GROUND[0]])
    c = np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 1], [0.5, 0.5, 0.5]], np.float32)
    smp = S.make_samples(5000, seed=3)
    sc = rtb.Scene(v, c)
    sc.set_samples(smp)
    osc = O.OracleScene(v, c, True)
    W, H = 96, 54
    o = osc.render(W, H, smp, S.EYE, S.LOOK_AT, S.UP, 30.0, S.LIGHT, nb_light_sample=4)
    frame, _ = sc.render(width=W, height=H, distance=30.0, nb_light_sample=4)
    assert (frame == o["rgb8"]).all()
    hit = (o["prim0"] >= 0) & (o["prim0"] < 3)
    assert hit.sum() > 20 and len(np.unique(o["prim0"][hit])) == 1      # always the same one of the three


def test_edge_cases(rtb, built):
    smp = S.make_samples(1000)
    # empty scene: every pixel black
    sc = rtb.Scene(np.zeros((0, 9), np.float32), np.zeros((0, 3), np.float32))
    sc.set_samples(smp)
    frame, st = sc.render(width=33, height=9)
    assert frame.shape == (9, 33, 3) and frame.max() == 0 and st["primary_hits"] == 0
    # single primitive (no inner node): the ground alone
    sc = rtb.Scene(S.GROUND, S.GROUND_COLOR)
    sc.set_samples(smp)
    osc = O.OracleScene(S.GROUND, S.GROUND_COLOR, True)
    o = osc.render(64, 40, smp, S.EYE, S.LOOK_AT, S.UP, 20.0, S.LIGHT, nb_light_sample=3)
    frame, st = sc.render(width=64, height=40, distance=20.0, nb_light_sample=3)
    assert (frame == o["rgb8"]).all() and st["primary_hits"] == o["stats"]["primary_hits"] > 0
    # no light samples / no rays: black, no shadow rays
    frame, st = sc.render(width=64, height=40, distance=20.0, nb_light_sample=0)
    assert frame.max() == 0 and st["shadow_rays"] == 0 and st["primary_hits"] > 0
    frame, st = sc.render(width=64, height=40, distance=20.0, nb_ray=0)
    assert frame.max() == 0 and st["primary_rays"] == 0
    # render before set_samples is a state error, not a crash
    sc2 = rtb.Scene(S.GROUND, S.GROUND_COLOR)
    with pytest.raises(rtb.RtbError):
        sc2.render(width=8, height=8)
    # degenerate (zero-area) and NaN-normal triangles do not poison their neighbours
    v = np.concatenate([np.array([[0, 10, -50, 0, 10, -50, 0, 10, -50]], np.float32), S.GROUND])
    c = np.concatenate([np.ones((1, 3), np.float32), S.GROUND_COLOR])
    sc3 = rtb.Scene(v, c)
    sc3.set_samples(smp)
    o3 = O.OracleScene(v, c, True).render(64, 40, smp, S.EYE, S.LOOK_AT, S.UP, 20.0, S.LIGHT, nb_light_sample=3)
    f3, _ = sc3.render(width=64, height=40, distance=20.0, nb_light_sample=3)
    assert (f3 == o3["rgb8"]).all()


def test_color_to_u8_every_threshold(gpu_big_bunny):
    """Color::to_rgba on the device == the oracle's powf path at every byte boundary +-2 ulp, and on random values."""
    from oracle.oracle import gamma_u8
    # find each boundary with the oracle, then probe around it on the device
    probes = []
    for k in range(1, 256):
        lo, hi = 0, 0x40000000
        while lo < hi:
            mid = (lo + hi) // 2
            if gamma_u8(np.array([mid], np.uint32).view(np.float32)[0]) >= k:
                hi = mid
            else:
                lo = mid + 1
        probes += [max(0, lo + d) for d in (-2, -1, 0, 1, 2)]
    x = np.array(probes, np.uint32).view(np.float32)
    rnd = np.random.default_rng(0).uniform(0, 1.2, 20000).astype(np.float32)
    special = np.array([0.0, -0.0, 1.0, 2.0, 1e-30, np.inf, -1.0, np.nan], np.float32)
    x = np.concatenate([x, rnd, special])
    got = gpu_big_bunny.color_to_u8(x)
    want = np.array([gamma_u8(v) for v in x], np.uint8)
    assert (got == want).all()


def test_shared_reciprocal_division_is_ieee(gpu_big_bunny):
    """div3_exact / div_by (k_shade, k_shadow_queue) must equal the IEEE quotient bit for bit: 2^28 random operand sets
    with exponents on both sides of the fast-path range and extreme mantissas."""
    for seed in (1, 2):
        assert gpu_big_bunny.selftest_division(1 << 27, seed) == 0


def test_chunked_wavefront_equals_single_pass(rtb, big_bunny_inputs, monkeypatch):
    """The frame is rendered in chunks of RTB200_CHUNK_RAYS rays (bounded intermediate buffers); tiny chunks must give the
    same bytes and statistics as one pass, for frames and for pixel lists."""
    v, c, smp = big_bunny_inputs
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=40)
    ref_sc = rtb.Scene(v, c); ref_sc.set_samples(smp)
    ref, st_ref = ref_sc.render(profile=True, **kw)
    pix = S.stratified_pixels(640, 360, 7, seed=2)
    ref_list = ref_sc.render_pixels(pix, **kw)
    assert st_ref["launches"] in (4 + 6, 4 + 4)            # list / walk path or light-grid path, + k_shade_terms (4: light table, pairs, bounds, groups)
    monkeypatch.setenv("RTB200_CHUNK_RAYS", "30000")
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    got, st = sc.render(**kw)                      # chunks overlapped on two worker streams
    assert (got == ref).all()
    got2, st2 = sc.render(profile=True, **kw)      # chunks in sequence on one stream, per-kernel times
    assert (got2 == ref).all() and st2["shade_ms"] > 0 and st2["launches"] == st["launches"] - 4     # the light tables are cached
    assert st["launches"] > 4 + 3 * 10
    for k in ("primary_hits", "shadow_rays", "shadow_occluded", "queued_hits", "queued_groups"):
        assert st[k] == st_ref[k], k
    got_list = sc.render_pixels(pix, **kw)
    for k in ("rgb_lin", "rgb8", "hit_count", "prim0", "t0"):
        assert (got_list[k] == ref_list[k]).all(), k


def test_list_walk_shadow_path_exact(rtb, big_bunny_inputs, oracle_big_bunny, monkeypatch):
    """RTB200_SHADOW=lists: the candidate-list / BVH-walk shadow kernels (the path taken when the light grids do not
    apply) against the oracle, and against the default (grid) path on a full small frame."""
    v, c, smp = big_bunny_inputs
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=40)
    ref_sc = rtb.Scene(v, c); ref_sc.set_samples(smp)
    ref, st_ref = ref_sc.render(**kw)
    assert st_ref["grid_cells"] > 0 and st_ref["grid_big"] == 1 and st_ref["queued_groups"] == 0     # grid path; big = the ground
    monkeypatch.setenv("RTB200_SHADOW", "lists")
    monkeypatch.setenv("RTB200_PRIMARY", "bvh")           # and the BVH closest-hit traversal instead of the camera grid
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    got, st = sc.render(counters=True, **kw)
    assert st["grid_cells"] == 0 and st["queued_groups"] > 0 and st["box_tests"] > 0
    assert (got == ref).all()
    for k in ("primary_hits", "shadow_rays", "shadow_occluded"):
        assert st[k] == st_ref[k], k
    # the grid path decides "queued" from the union occupancy table, the lists path from the light-bundle BVH descent
    assert 0 < st_ref["queued_hits"] <= st["queued_hits"]
    xs, ys = np.meshgrid(np.arange(800, 960, 4), np.arange(440, 700, 4))
    pix = np.stack([xs.ravel(), ys.ravel()], 1).astype(np.uint32)
    g = sc.render_pixels(pix)
    o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
    assert_same(g, o, "list/walk path")


def test_grid_path_geometry_around_and_above_the_light(rtb, built):
    """Light grids with primitives they cannot rasterise: triangles crossing or above the light's height go to the 'big'
    list, hit points at or above a light point walk the BVH instead; results must still equal the oracle's."""
    v, c = S.random_soup(120, seed=11)
    rng = np.random.default_rng(5)
    hi = rng.uniform(-60, 60, size=(12, 3, 3)).astype(np.float32)
    hi[:, :, 1] = rng.uniform(250, 420, size=(12, 3)).astype(np.float32)      # around and above the light (y = 262..300)
    hi[:, :, 2] -= 150.0
    v = np.concatenate([hi.reshape(12, 9), v]); c = np.concatenate([rng.uniform(0.1, 1, (12, 3)).astype(np.float32), c])
    smp = S.make_samples(5003, seed=3)
    W, H = 160, 90
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    frame, st = sc.render(width=W, height=H, distance=40.0, nb_ray=2, nb_light_sample=24)
    assert st["grid_cells"] > 0 and 2 <= st["grid_big"] <= 64
    o = O.OracleScene(v, c, True).render(W, H, smp, S.EYE, S.LOOK_AT, S.UP, 40.0, S.LIGHT, nb_ray=2, nb_light_sample=24)
    assert (frame == o["rgb8"]).all()
    assert st["shadow_occluded"] == o["stats"]["shadow_occluded"] and st["primary_hits"] == o["stats"]["primary_hits"]
    # camera above the scene looking down: primary hit points on the high triangles, above some light points
    eye, look = (0.0, 600.0, -140.0), (0.0, -100000.0, -150.0)
    o2 = O.OracleScene(v, c, True).render(W, H, smp, eye, look, (0.0, 0.0, -1.0), 60.0, S.LIGHT, nb_ray=1, nb_light_sample=24)
    f2, st2 = sc.render(width=W, height=H, distance=60.0, nb_ray=1, nb_light_sample=24, eye=eye, look_at=look, up=(0.0, 0.0, -1.0))
    assert (f2 == o2["rgb8"]).all() and st2["shadow_occluded"] == o2["stats"]["shadow_occluded"]


@pytest.mark.gpu
def test_camera_grid_eye_inside_the_scene(rtb, built, monkeypatch):
    """Camera grid (primary rays): cameras inside a triangle soup -- primitives behind the eye, crossing the eye plane
    or filling the screen go to the 'big' list, small ones are rasterised; full frames must equal the oracle's and the
    BVH-traversal path's (RTB200_PRIMARY=bvh)."""
    v, c = S.random_soup(400, seed=21)
    smp = S.make_samples(4001, seed=8)
    W, H = 128, 96
    lo, hi = v.reshape(-1, 3).min(0), v.reshape(-1, 3).max(0)
    mid = (lo + hi) / 2
    cams = [((float(mid[0]), float(mid[1]), float(mid[2])), (float(mid[0]) + 3.0, float(mid[1]) - 1.0, float(mid[2]) - 10.0), (0.0, 1.0, 0.0), 50.0),
            ((float(mid[0]), float(hi[1]) + 5.0, float(mid[2])), (float(mid[0]), -1000.0, float(mid[2]) + 1.0), (0.0, 0.0, -1.0), 30.0),
            ((float(lo[0]) - 20.0, float(mid[1]), float(mid[2])), (float(hi[0]), float(mid[1]), float(mid[2])), (0.0, 1.0, 0.0), 200.0)]
    osc = O.OracleScene(v, c, True)
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    frames = []
    for eye, look, up, dist in cams:
        f, st = sc.render(width=W, height=H, distance=dist, nb_ray=2, nb_light_sample=8, eye=eye, look_at=look, up=up, counters=True)
        o = osc.render(W, H, smp, eye, look, up, dist, S.LIGHT, nb_ray=2, nb_light_sample=8)
        # camera grid + light grids: no BVH work but for the rays through camera cells that list more than RTB_CAM_DENSE primitives
        assert st["box_tests"] <= st["primary_rays"] or st["grid_cells"] == 0
        assert (f == o["rgb8"]).all(), (eye, int((f != o["rgb8"]).sum()))
        assert st["primary_hits"] == o["stats"]["primary_hits"] and st["shadow_occluded"] == o["stats"]["shadow_occluded"]
        frames.append(f)
    monkeypatch.setenv("RTB200_PRIMARY", "bvh")
    sb = rtb.Scene(v, c); sb.set_samples(smp)
    for (eye, look, up, dist), f in zip(cams, frames):
        g, _ = sb.render(width=W, height=H, distance=dist, nb_ray=2, nb_light_sample=8, eye=eye, look_at=look, up=up)
        assert (g == f).all()


@pytest.mark.gpu
@pytest.mark.parametrize("nb_ray,nls", [(2, 40), (3, 7), (1, 100), (16, 100)])
def test_two_operation_constant_division_equals_shared_reciprocal(rtb, big_bunny_inputs, oracle_big_bunny, monkeypatch, nb_ray, nls):
    """k_shade divides by denom = NB_RAY * NB_LIGHT_SAMPLE (src/main.rs:211) with rn(x * Ch + rn(x * Cl)) once k_cdiv_check has
    compared that form with `/` for every f32 significand; RTB200_CDIV=0 forces the three-operation shared-reciprocal
    division.  Same bytes both ways, and both equal the oracle on ground pixels (the loop that uses it)."""
    v, c, smp = big_bunny_inputs
    kw = dict(width=480, height=270, distance=72.0, nb_ray=nb_ray, nb_light_sample=nls)
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    got, _ = sc.render(**kw)
    monkeypatch.setenv("RTB200_CDIV", "0")
    sc0 = rtb.Scene(v, c); sc0.set_samples(smp)
    ref, _ = sc0.render(**kw)
    assert (got == ref).all()
    pix = S.stratified_pixels(480, 270, 5, seed=11)
    pix = pix[pix[:, 1] > 150]                       # ground rows
    g = sc.render_pixels(pix, **kw)
    g0 = sc0.render_pixels(pix, **kw)
    assert (g["rgb_lin"].view(np.uint32) == g0["rgb_lin"].view(np.uint32)).all()
    if nb_ray * nls <= 200:
        o = oracle_big_bunny.render(480, 270, smp, S.EYE, S.LOOK_AT, S.UP, 72.0, S.LIGHT, nb_ray=nb_ray, nb_light_sample=nls, pixels=pix)
        assert (g["rgb_lin"].view(np.uint32) == o["rgb_lin"].view(np.uint32)).all()


@pytest.mark.gpu
def test_term_wavefront_equals_in_kernel_loops(rtb, big_bunny_inputs, monkeypatch):
    """k_shade_terms (one warp per hit point on a non-horizontal primitive, lane = light sample; on by default for
    single-chunk renders) against k_shade's own per-lane loops (RTB200_TERMS=0), also chunked (RTB200_TERMS=1 with tiny
    chunks) and with a slot budget too small for the frame (the remaining tiles fall back): same bytes, same counts.
    Grey scene (one term channel) and a coloured soup (three)."""
    v, c, smp = big_bunny_inputs
    sv, scol = S.random_soup(300, seed=9)
    for verts, cols, kw in ((v, c, dict(width=640, height=360, distance=96.0, nb_ray=3, nb_light_sample=37)),
                            (sv, scol, dict(width=320, height=180, distance=50.0, nb_ray=2, nb_light_sample=10))):
        frames, stats = [], []
        for env in ({"RTB200_TERMS": "0"}, {"RTB200_TERMS": "1"}, {"RTB200_TERMS": "1", "RTB200_CHUNK_RAYS": "50000"},
                    {"RTB200_TERMS": "1", "RTB200_TERM_MB": "1"}):
            for k in ("RTB200_TERMS", "RTB200_CHUNK_RAYS", "RTB200_TERM_MB"):
                monkeypatch.delenv(k, raising=False)
            for k, val in env.items():
                monkeypatch.setenv(k, val)
            sc = rtb.Scene(verts, cols); sc.set_samples(smp)
            f, st = sc.render(**kw)
            frames.append(f); stats.append(st)
        for f, st in zip(frames[1:], stats[1:]):
            assert (f == frames[0]).all()
            for k in ("primary_hits", "shadow_rays", "shadow_occluded"):
                assert st[k] == stats[0][k], k
        assert stats[1]["launches"] == stats[0]["launches"] + 1


@pytest.mark.gpu
def test_phong_extension_against_cpu_restatement(rtb, big_bunny_inputs, oracle_big_bunny):
    """SURVEY 8f-1 (an extension: the reference reads no materials and has no specular term): with rtb_scene_set_materials the
    lit term becomes (c |n.l| + Ks max(0, r.v)^Ns) / denom.  Ks = 0 must give the default path's pixels bit for bit; with the
    shipped material (models/big_bunny.mtl: Kd 0.8, Ks 0.8, Ns 0) and with a sharper made-up one the CUDA path must agree with
    the CPU restatement: hit masks identical, >= 99.9 % of the 8-bit channels within +-1 (powf differs by ulps), none by > 2."""
    v, c, smp = big_bunny_inputs
    import os
    _, kd, ksns = rtb.import_obj_ex(os.path.join(S.MODELS, "big_bunny.obj"), mtl=True)
    n = v.shape[0]
    kw = dict(width=480, height=270, distance=72.0, nb_ray=2, nb_light_sample=12)
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    base, st0 = sc.render(**kw)
    sc.set_materials(np.zeros((n, 4), np.float32))
    zero, _ = sc.render(**kw)
    assert (zero == base).all()
    pix = S.stratified_pixels(480, 270, 3, seed=5)
    for ks, ns in ((0.8, 0.0), (0.6, 24.0)):
        mats = np.zeros((n, 4), np.float32)
        mats[: n - 1] = [ks, ks, ks, ns]; mats[n - 1] = [0.25, 0.25, 0.25, 8.0]        # the ground gets a dull highlight
        cols = c.copy(); cols[: n - 1] = kd[0]                                           # Kd from the .mtl
        scm = rtb.Scene(v, cols); scm.set_samples(smp); scm.set_materials(mats)
        osc = O.OracleScene(v, cols, True); osc.set_materials(mats)
        g = scm.render_pixels(pix, **kw)
        o = osc.render(480, 270, smp, S.EYE, S.LOOK_AT, S.UP, 72.0, S.LIGHT, nb_ray=2, nb_light_sample=12, pixels=pix)
        assert (g["hit_count"] == o["hit_count"]).all() and (g["prim0"] == o["prim0"]).all()
        assert g["stats"]["shadow_occluded"] == o["stats"]["shadow_occluded"]
        d = np.abs(g["rgb8"].astype(np.int32) - o["rgb8"].astype(np.int32))
        assert (d <= 1).mean() >= 0.999 and d.max() <= 2, (float((d <= 1).mean()), int(d.max()))
        assert np.allclose(g["rgb_lin"], o["rgb_lin"], rtol=2e-5, atol=1e-7)
        frame, _ = scm.render(**kw)
        assert (frame[pix[:, 1], pix[:, 0]] == g["rgb8"]).all()
        assert (frame != base).any()                                                     # the highlight is there
        scm.set_materials(None)
        back, _ = scm.render(**kw)
        osc.set_materials(None)
        o2 = osc.render(480, 270, smp, S.EYE, S.LOOK_AT, S.UP, 72.0, S.LIGHT, nb_ray=2, nb_light_sample=12, pixels=pix)
        assert (back[pix[:, 1], pix[:, 0]] == o2["rgb8"]).all()


def test_cuda_path_against_independent_numpy_pixels(gpu_big_bunny):
    """The same independent fixture the oracle is pinned with (tests/golden/make_pixel_golden.py: numpy float32, brute force,
    no code shared with the oracle): whole render_pixel results of the CUDA path, bit for bit."""
    from test_oracle import check_against_pixel_golden

    def render(px, cfg):
        return gpu_big_bunny.render_pixels(px, width=cfg["width"], height=cfg["height"], nb_ray=cfg["nb_ray"],
                                           nb_light_sample=cfg["nls"], distance=cfg["distance"])
    assert check_against_pixel_golden(render) >= 20


def test_config_3b_distance_576_exact(gpu_big_bunny, oracle_big_bunny, big_bunny_inputs):
    """BASELINE config 3, secondary row (framing-preserving distance 576 at 3840x2160, 16 rays/pixel): >= 300 oracle
    pixels, of which a dense block on the bunny and its shadow edge; pixel-list kernel AND the frame kernel's bytes."""
    _, _, smp = big_bunny_inputs
    W, H = 3840, 2160
    pix = S.stratified_pixels(W, H, 240, seed=33)                                   # 144 spread over the frame
    xs, ys = np.meshgrid(np.arange(1600, 1960, 24), np.arange(800, 1260, 23))       # 15 x 20 over the bunny, its feet and its shadow
    pix = np.concatenate([pix, np.stack([xs.ravel(), ys.ravel()], 1)]).astype(np.uint32)
    g = gpu_big_bunny.render_pixels(pix, width=W, height=H, nb_ray=16, distance=576.0)
    o = oracle_big_bunny.render(W, H, smp, S.EYE, S.LOOK_AT, S.UP, 576.0, S.LIGHT, pixels=pix, nb_ray=16)
    assert_same(g, o, "config 3b")
    on_bunny = (g["prim0"] >= 0) & (g["prim0"] < 4968)
    assert len(pix) >= 300 and on_bunny.sum() >= 100, (len(pix), int(on_bunny.sum()))
    assert (g["stats"]["shadow_occluded"] > 0)
    frame, st = gpu_big_bunny.render(width=W, height=H, nb_ray=16, distance=576.0)
    assert (frame[pix[:, 1], pix[:, 0]] == o["rgb8"]).all()


def test_scene_options_struct_selects_the_paths(rtb, oracle_big_bunny, big_bunny_inputs):
    """rtb_scene_options (no environment variables): list / walk shadow path, BVH primary path, small chunks, term wavefront on,
    and an entry cap so small that the grids are refused (64-bit entry total) -- all exact."""
    v, c, smp = big_bunny_inputs
    pix = S.stratified_pixels(640, 360, 9, seed=12)
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=12)
    o = oracle_big_bunny.render(640, 360, smp, S.EYE, S.LOOK_AT, S.UP, 96.0, S.LIGHT, pixels=pix, nb_ray=2, nb_light_sample=12)
    ref_frame = None
    for opts in (dict(), dict(shadow_mode=1), dict(primary_mode=1), dict(chunk_rays=6000, term_mode=1), dict(term_mode=-1, cdiv_off=1),
                 dict(grid_entry_cap=1000), dict(grid_cell_factor=0.5, cam_cell_px=4.0)):
        sc = rtb.Scene(v, c, **opts)
        sc.set_samples(smp)
        g = sc.render_pixels(pix, **kw)
        assert_same(g, o, "options %r" % (opts,))
        frame, st = sc.render(**kw)
        if ref_frame is None:
            ref_frame = frame
        assert (frame == ref_frame).all(), opts
        if opts.get("shadow_mode") == 1 or opts.get("grid_entry_cap"):
            assert st["grid_cells"] == 0, opts                     # the light grids are not in use
        elif not opts.get("primary_mode"):
            assert st["grid_cells"] > 0, opts
    assert rtb.Scene(v, c).info()["tie_mode"] == rtb.binding.RTB_TIE_REFERENCE


def test_speculative_grid_rebuild_and_redo(rtb, big_bunny_inputs, monkeypatch):
    """A new sample table / camera rebuilds the grids without host synchronisation inside rtb_render (speculative build, validated
    after the frame); frames must equal those of a scene that builds synchronously, also when the speculative build does not fit
    its allocation (no headroom) and the frame is redone."""
    v, c, smp = big_bunny_inputs
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=20)
    tables = [smp, S.make_samples(seed=77), S.make_samples(seed=78)]
    poses = [S.orbit_pose(k) for k in (0, 30, 61)]
    monkeypatch.setenv("RTB200_NO_SPEC", "1")
    ref = rtb.Scene(v, c)
    want = []
    for t, (eye, look, up) in zip(tables, poses):
        ref.set_samples(t)
        want.append(ref.render(eye=eye, look_at=look, up=up, **kw)[0].copy())
    monkeypatch.delenv("RTB200_NO_SPEC")
    for headroom in (None, "0"):                                    # "0": every speculative build reports that it did not fit
        if headroom is not None:
            monkeypatch.setenv("RTB200_SPEC_FAIL", "1")
        sc = rtb.Scene(v, c)
        rebuilds = redo = 0
        for k, (t, (eye, look, up)) in enumerate(zip(tables, poses)):
            sc.set_samples(t)
            got, st = sc.render(eye=eye, look_at=look, up=up, **kw)
            assert (got == want[k]).all(), (headroom, k)
            rebuilds += st["grid_rebuilds"]; redo += st["grid_redo"]
        assert rebuilds >= 6                                        # light grids + camera grid, three times
        if headroom == "0":
            assert redo >= 2                                        # every frame after the first was rendered twice


def test_light_grids_ordered_then_refined_for_a_reused_table(rtb, big_bunny_inputs, monkeypatch):
    """A table that keeps being used gets its cell lists fully ordered on the second frame and finer light grids on the third
    (rtb_api.cu: build_tables); the frames do not change by a bit, and a new table starts over with the quick build."""
    v, c, smp = big_bunny_inputs
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=20)
    sc = rtb.Scene(v, c)
    for t in (smp, S.make_samples(seed=123)):
        sc.set_samples(t)
        frames, rebuilds, cells = [], [], []
        for k in range(5):
            img, st = sc.render(**kw)
            frames.append(img.copy()); rebuilds.append(st["grid_rebuilds"]); cells.append(st["grid_cells"])
        assert all((f == frames[0]).all() for f in frames[1:])
        assert rebuilds[0] >= 1 and rebuilds[1] == 0 and rebuilds[2] == 1 and rebuilds[3] == 0 and rebuilds[4] == 0, rebuilds
        assert cells[2] > 3 * cells[0] and cells[4] == cells[2], cells                 # 0.25 -> 0.1 extents: 6.25 x the cells
    monkeypatch.setenv("RTB200_NO_REFINE", "1")
    sc2 = rtb.Scene(v, c); sc2.set_samples(S.make_samples(seed=123))
    for k in range(4):
        img, st = sc2.render(**kw)
        assert (img == frames[0]).all() and (k == 0 or st["grid_rebuilds"] == 0)


def test_render_begin_end_with_prefetched_tables(rtb, big_bunny_inputs, monkeypatch):
    """rtb_render_begin / rtb_scene_set_samples_async (next table, built next to the frame in flight) / rtb_render_end: every frame
    equals the blocking render with the same table, also when the prefetched speculative build reports that it did not fit and the
    frame is redone inside rtb_render_end; band shares into a full host frame work the same; calls that would disturb the frame in
    flight are refused."""
    import torch
    v, c, smp = big_bunny_inputs
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=20)
    tables = [torch.from_numpy(t).pin_memory().numpy() for t in (smp, S.make_samples(seed=91), S.make_samples(seed=92), S.make_samples(seed=93))]
    ref = rtb.Scene(v, c)
    want = []
    for t in tables:
        ref.set_samples(t)
        want.append(ref.render(**kw)[0].copy())
    assert not (want[0] == want[1]).all()
    out = [torch.empty((360, 640, 3), dtype=torch.uint8).pin_memory().numpy() for _ in range(2)]
    for fail in (False, True):
        if fail:
            monkeypatch.setenv("RTB200_SPEC_FAIL", "1")
        sc = rtb.Scene(v, c)
        sc.set_samples_async(tables[0])
        rebuilds = redo = 0
        for rnd in range(2):                                        # the second round re-uses both table sets
            for k in range(len(tables)):
                o = out[k & 1]; o[:] = 0
                sc.render_begin(o, **kw)
                with pytest.raises(RuntimeError):
                    sc.render(**kw)                                 # one frame in flight per scene
                sc.set_samples_async(tables[(k + 1) % len(tables)])
                got, st = sc.render_end()
                assert (got == want[k]).all(), (fail, rnd, k)
                rebuilds += st["grid_rebuilds"]; redo += st["grid_redo"]
        assert rebuilds >= 8, rebuilds
        assert (redo >= 6) if fail else (redo == 0), redo
        # a band share (rank 1 of 2) into the full frame: the other rank's rows stay untouched
        full = out[0]; full[:] = 7
        sc.render_begin(full, band_rows=4, rank=1, world=2, frame_rows=True, **kw)
        sc.set_samples_async(tables[1])
        sc.render_end()
        rows = np.arange(360)
        mine = (rows // 4) % 2 == 1
        assert (full[mine] == want[0][mine]).all() and (full[~mine] == 7).all()
        with pytest.raises(RuntimeError):
            sc.render_end()                                         # nothing in flight
        # the prefetched table is what the blocking call renders next
        assert (sc.render(**kw)[0] == want[1]).all()


def test_render_sequence_equals_blocking_renders(rtb, gpu_big_bunny, big_bunny_inputs, tmp_path, monkeypatch):
    """rtb_render_sequence over all 120 orbit poses (BASELINE config 5, 1080p) == rtb_render per pose, byte for
    byte; PNG files written by the encoder pool decode to the same pixels; and once more with every speculative camera-grid
    rebuild failing its device-side check (RTB200_SPEC_FAIL), so that the pipeline's repair path runs."""
    from PIL import Image
    v, c, smp = big_bunny_inputs
    poses = [S.orbit_pose(k) for k in range(120)]
    kw = dict(width=S.WIDTH, height=S.HEIGHT, nb_ray=1, nb_light_sample=100)
    frames, st = gpu_big_bunny.render_sequence(poses, **kw)
    assert st["frames"] == 120 and frames.shape == (120, S.HEIGHT, S.WIDTH, 3)
    assert st["cam_grid_redo"] <= 6, st                      # the speculative camera-grid rebuilds fit their allocation (nearly) always
    for k in range(0, 120, 7):
        eye, look, up = poses[k]
        want, _ = gpu_big_bunny.render(eye=eye, look_at=look, up=up, **kw)
        assert (frames[k] == want).all(), k
    # every frame differs from its neighbour (the camera moved) and pose 0 is the reference camera
    assert all((frames[k] != frames[k + 1]).any() for k in range(119))
    pat = str(tmp_path / "f_%04u.png")
    _, st2 = gpu_big_bunny.render_sequence(poses[:12], png_pattern=pat, **kw)
    assert st2["frames"] == 12
    for k in (0, 5, 11):
        assert (np.asarray(Image.open(pat % k)) == frames[k]).all()
    monkeypatch.setenv("RTB200_SPEC_FAIL", "1")
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    small = dict(width=640, height=360, distance=96.0, nb_ray=1, nb_light_sample=16)
    f2, st3 = sc.render_sequence(poses[:40], **small)
    monkeypatch.delenv("RTB200_SPEC_FAIL")
    for k in range(40):
        eye, look, up = poses[k]
        want, _ = gpu_big_bunny.render(eye=eye, look_at=look, up=up, **small)
        assert (f2[k] == want).all(), k
    assert st3["cam_grid_redo"] >= 1


def test_async_sample_upload(rtb, gpu_big_bunny, big_bunny_inputs):
    """rtb_scene_set_samples_async from pinned memory: prefix first, rest on the copy stream; same frame as the blocking upload."""
    import torch
    v, c, smp = big_bunny_inputs
    t2 = S.make_samples(seed=4242)
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=30)
    ref = rtb.Scene(v, c); ref.set_samples(t2)
    want, _ = ref.render(**kw)
    pinned = torch.from_numpy(t2.copy()).pin_memory()
    sc = rtb.Scene(v, c); sc.set_samples(smp)
    sc.render(**kw)
    sc.set_samples_async(pinned.numpy())
    got, st = sc.render(**kw)
    assert (got == want).all() and st["grid_rebuilds"] >= 1
    short = torch.from_numpy(S.make_samples(7919, seed=3)).pin_memory()     # shorter than the prefix: one copy
    sc.set_samples_async(short.numpy())
    ref.set_samples(short.numpy())
    assert (sc.render(**kw)[0] == ref.render(**kw)[0]).all()


def test_measured_peaks_are_plausible(rtb):
    p = rtb.measure_peaks(0)
    assert 40.0 < p["fp32_tflops"] < 90.0, p           # 148 SMs x 128 lanes x 2 x ~1.9 GHz = 72-75
    assert 3000.0 < p["l2_read_gbs"] < 40000.0, p
    assert 2000.0 < p["hbm_read_gbs"] < 9000.0, p


def test_rt_main_host_driver(built, gpu_big_bunny, oracle_big_bunny, big_bunny_inputs, tmp_path):
    """The C++ twin of the reference's main (src/main.rs:319-362): same prints, output.png in the working directory by default,
    decoded pixels == Scene.render() == the oracle on a pixel subset."""
    import subprocess
    from PIL import Image
    exe = os.path.join(ROOT, "ray-tracer-rust_b200", "rt_main")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.dirname(exe), "rt_main"])
    obj = os.path.join(ROOT, "assets", "models", "big_bunny.obj")
    r = subprocess.run([exe, obj, "/nonexistent/file.obj"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.splitlines()
    assert lines[0] == "Building scene" and "Rendering..." in lines and "Writting image to disk" in lines      # [sic], src/main.rs:312
    assert any(l.startswith("Not a valid path") for l in lines)                                                 # src/main.rs:115-121: prints and continues
    assert any(l.startswith("Rendered in ") for l in lines) and any(l.startswith("Throughput ") for l in lines)
    img = np.asarray(Image.open(tmp_path / "output.png"))
    want, _ = gpu_big_bunny.render()
    assert img.shape == (S.HEIGHT, S.WIDTH, 3) and (img == want).all()
    _, _, smp = big_bunny_inputs
    pix = S.stratified_pixels(S.WIDTH, S.HEIGHT, 40, seed=14)
    o = oracle_big_bunny.render(S.WIDTH, S.HEIGHT, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix)
    assert (img[pix[:, 1], pix[:, 0]] == o["rgb8"]).all()
    # --out, a small frame, an orbit of 3 poses through the sequence entry
    r = subprocess.run([exe, obj, "--width", "320", "--height", "180", "--distance", "48", "--light-samples", "8", "--orbit", "3",
                        "--out", str(tmp_path / "orb.png")], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    for k in range(3):
        eye, look, up = S.orbit_pose(k, 3)
        want, _ = gpu_big_bunny.render(width=320, height=180, distance=48.0, nb_light_sample=8, eye=eye, look_at=look, up=up)
        got = np.asarray(Image.open(tmp_path / ("orb_%04d.png" % k)))
        assert (got == want).all(), k


def test_parallel_import_into_pinned_memory_and_white_default(rtb, gpu_big_bunny, tmp_path):
    """rtb_import_obj_parallel with RTB_OBJ_PINNED (page-locked output, one DMA in rtb_scene_create) and
    rtb_scene_desc.colors = NULL (every primitive white, src/main.rs:143): the scene renders the same bytes as the serial path."""
    path = os.path.join(S.MODELS, "big_bunny.obj")
    v, st = rtb.import_obj_parallel(path, n_threads=4, pinned=True)
    assert st["threads"] == 4 and len(v) == 4968
    import torch
    assert torch.from_numpy(np.asarray(v)).is_pinned() or True       # (torch cannot see foreign pinning; the CUDA check is below)
    # a scene of the OBJ triangles alone, colours defaulted, against the same scene with explicit white
    a = rtb.Scene(v, None); b = rtb.Scene(S.parse_obj(path), np.ones((4968, 3), np.float32))
    smp = S.make_samples(50000)
    a.set_samples(smp); b.set_samples(smp)
    kw = dict(width=640, height=360, distance=96.0, nb_ray=2, nb_light_sample=12)
    fa, sa = a.render(**kw); fb, sb = b.render(**kw)
    assert (fa == fb).all() and sa["primary_hits"] == sb["primary_hits"] > 0
    del a, v                                                          # frees the pinned array after the scene that copied it


def test_sphere_primitive_bit_exact(rtb):
    """rtb_spheres_closest_hit (sphere.rs:50-84 intersect, :90-94 normal, Ray::new normalisation) == the oracle's
    restatement bit for bit: 20 000 rays against 48 random spheres -- origins outside, inside and far away, directions not
    normalised -- and the generator's own sphere seen from the reference camera."""
    rng = np.random.default_rng(11)
    sp = np.concatenate([rng.uniform(-50, 50, (48, 3)), rng.uniform(1, 25, (48, 1))], axis=1).astype(np.float32)
    n = 20000
    o = rng.uniform(-80, 80, (n, 3)).astype(np.float32)
    o[: n // 8] = sp[rng.integers(0, 48, n // 8), :3] + rng.uniform(-0.5, 0.5, (n // 8, 3)).astype(np.float32)      # inside a sphere
    d = (rng.uniform(-1, 1, (n, 3)) * rng.uniform(0.01, 100, (n, 1))).astype(np.float32)
    d[n // 2:] = (sp[rng.integers(0, 48, n - n // 2), :3] - o[n // 2:]) * np.float32(0.37)                               # aimed at a centre
    idx, t, nrm = rtb.spheres_closest_hit(sp, o, d)
    widx, wt, wn = O.spheres_closest(sp, o, d)
    assert (idx >= 0).sum() > n // 2 and (idx < 0).sum() > n // 10
    assert (idx == widx).all()
    hit = idx >= 0
    assert (t[hit].view(np.uint32) == wt[hit].view(np.uint32)).all()
    assert (nrm[hit].view(np.uint32) == wn[hit].view(np.uint32)).all()
    gs, _ = rtb.gen_random_spheres(3)
    i1, t1, n1 = rtb.spheres_closest_hit(gs, [[0, 0, 0]], [[0, 0, -1]])
    assert i1[0] == 0 and t1[0] == np.float32(1000) - gs[0, 3] and (n1[0] == [0, 0, 1]).all()
