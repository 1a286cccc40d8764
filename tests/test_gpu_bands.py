"""GPU tests of the multi-GPU decomposition on ONE device: every rank's band set rendered in turn must tile the
single-GPU frame byte for byte, and the de-interleave kernel must equal the host bookkeeping."""
import importlib

import numpy as np
import pytest

import scenes as S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scene(built, big_bunny_inputs):
    import rtb200
    v, c, smp = big_bunny_inputs
    sc = rtb200.Scene(v, c)
    sc.set_samples(smp)
    return sc


@pytest.mark.parametrize("world,band_rows,w,h", [(2, 8, 640, 360), (3, 8, 333, 187), (8, 8, 640, 360), (4, 16, 640, 360), (8, 1, 64, 37)])
def test_bands_tile_the_frame(scene, world, band_rows, w, h):
    import rtb200
    D = importlib.import_module("ray_tracer_rust_b200.distributed")
    kw = dict(width=w, height=h, distance=288.0 * w / 1920, nb_light_sample=8)
    full, st_full = scene.render(**kw)
    parts, hits = [], 0
    for r in range(world):
        rows, st = scene.render(band_rows=band_rows, rank=r, world=world, **kw)
        assert rows.shape[0] == rtb200.owned_rows(h, band_rows, r, world)
        parts.append(rows)
        hits += st["primary_hits"]
    assert hits == st_full["primary_hits"]
    assert (D.assemble_on_host(parts, w, h, band_rows, world) == full).all()


def test_deinterleave_kernel(scene):
    import torch
    import rtb200
    D = importlib.import_module("ray_tracer_rust_b200.distributed")
    for (w, h, band, world) in ((640, 360, 8, 4), (333, 187, 8, 3), (64, 37, 1, 8)):
        stride = D.padded_stride(w, h, band, world)
        rng = np.random.default_rng(w)
        staged = rng.integers(0, 256, (world, stride), dtype=np.uint8)
        d_staged = torch.from_numpy(staged).cuda()
        d_frame = torch.zeros(h * w * 3, dtype=torch.uint8, device="cuda")
        rtb200.deinterleave_bands(0, d_staged.data_ptr(), stride, d_frame.data_ptr(), w, h, band, world,
                                  torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        want = D.assemble_on_host([staged[r] for r in range(world)], w, h, band, world)
        assert (d_frame.cpu().numpy().reshape(h, w, 3) == want).all()


def test_render_device_into_torch_buffer(scene):
    """The NCCL-send-buffer form: asynchronous render into caller-owned device memory on the caller's stream."""
    import torch
    D = importlib.import_module("ray_tracer_rust_b200.distributed")
    w, h = 640, 360
    g = D.BandGather(w, h, 8, 0, 1, "cuda:0")
    frame = D.render_distributed(scene, g, distance=96.0, nb_light_sample=8)
    torch.cuda.synchronize()
    want, _ = scene.render(width=w, height=h, distance=96.0, nb_light_sample=8)
    assert (frame.cpu().numpy().reshape(h, w, 3) == want).all()


def _peer_worker(rank, world, port, w, h, band, q):
    import os, sys
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import rtb200
    import scenes as S2
    D = importlib.import_module("ray_tracer_rust_b200.distributed")
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)      # both ranks share cuda:0 here, which NCCL refuses
    try:
        v, c = S2.big_bunny()
        sc = rtb200.Scene(v, c); sc.set_samples(S2.make_samples())
        kw = dict(distance=288.0 * w / 1920, nb_ray=2, nb_light_sample=8)
        pf = D.PeerFrame(w, h, band, rank, world, "cuda:0")
        if rank == 0:
            for f in pf.frames:
                f.zero_()
        torch.cuda.synchronize(); dist.barrier()
        for _ in range(2):                                            # twice: the mappings are reused frame after frame
            frame = D.render_distributed(sc, pf, **kw)
        torch.cuda.synchronize(); dist.barrier()
        first = frame.cpu().numpy().reshape(h, w, 3).copy() if rank == 0 else None
        # two different poses back to back without waiting in between: frame A must survive the launch of frame B
        # (a single buffer could show rows of frame N + 1 to a reader of frame N)
        eyeB, lookB, upB = S2.orbit_pose(40)
        frameA = D.render_distributed(sc, pf, **kw)
        frameB = D.render_distributed(sc, pf, eye=eyeB, look_at=lookB, up=upB, **kw)
        torch.cuda.synchronize(); dist.barrier()
        if rank == 0:
            full, _ = sc.render(width=w, height=h, **kw)
            fullB, _ = sc.render(width=w, height=h, eye=eyeB, look_at=lookB, up=upB, **kw)
            ok = bool((first == full).all()) and bool((frameA.cpu().numpy().reshape(h, w, 3) == full).all()) \
                and bool((frameB.cpu().numpy().reshape(h, w, 3) == fullB).all()) and frameA.data_ptr() != frameB.data_ptr()
            q.put(ok)
        pf.close()
    finally:
        dist.destroy_process_group()


def test_peer_frame_two_processes_one_kernel_gather(built):
    """Render + gather as one kernel: two processes (both on cuda:0), rank 1 maps rank 0's frame through CUDA IPC and its
    shading kernel writes its bands straight into it; the result must equal the single-process frame."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    w, h, band, world = 320, 180, 8, 2
    procs = [ctx.Process(target=_peer_worker, args=(r, world, 29631, w, h, band, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert q.get(timeout=5) is True


def test_render_into_pinned_host_rows_chunk_by_chunk(built, big_bunny_inputs, monkeypatch):
    """rtb_render with a pinned destination copies each chunk's rows behind its k_shade on the chunk's own stream; same
    bytes as the single copy into pageable memory, for a full frame and for a band share, with many small chunks."""
    import torch
    import rtb200
    v, c, smp = big_bunny_inputs
    monkeypatch.setenv("RTB200_CHUNK_RAYS", "6000")
    sc = rtb200.Scene(v, c); sc.set_samples(smp)
    for kw in (dict(width=640, height=362, distance=96.0, nb_ray=2, nb_light_sample=9),
               dict(width=333, height=187, distance=60.0, nb_ray=1, nb_light_sample=5, band_rows=4, rank=1, world=3)):
        want, st = sc.render(**kw)
        assert st["launches"] > 8                                    # several chunks
        pinned = torch.zeros(want.size, dtype=torch.uint8).pin_memory()
        got, _ = sc.render(out=pinned.numpy().reshape(want.shape), **kw)
        assert (got == want).all()


def test_full_size_4k_16spp_band_shares_tile_the_frame(scene):
    """BASELINE config 3 at its full size (3840x2160, 16 rays/pixel, 100 light samples): the 8 band shares of the bench
    (4-row bands; each a single chunk, so the term wavefront runs) must tile the full frame (8 chunks on two streams, k_shade's
    own loops) byte for byte, with the same ray counts -- a size-independent property; the oracle comparison of this config
    is test_gpu_parity.py::test_4k_16spp_subset_exact."""
    D = importlib.import_module("ray_tracer_rust_b200.distributed")
    kw = dict(width=3840, height=2160, nb_ray=16, nb_light_sample=100)
    full, st_full = scene.render(**kw)
    assert st_full["launches"] >= 8 * 3
    parts, sums = [], {"primary_hits": 0, "shadow_rays": 0, "shadow_occluded": 0}
    for r in range(8):
        rows, st = scene.render(band_rows=4, rank=r, world=8, **kw)
        parts.append(rows)
        for k in sums:
            sums[k] += st[k]
    for k in sums:
        assert sums[k] == st_full[k], k
    assert (D.assemble_on_host(parts, 3840, 2160, 4, 8) == full).all()
    # pixels of the frame against the per-pixel entry point (the one the oracle tests use)
    pix = S.stratified_pixels(3840, 2160, 97, seed=21)
    g = scene.render_pixels(pix, **kw)
    assert (full[pix[:, 1], pix[:, 0]] == g["rgb8"]).all()


def test_full_size_4k_16spp_frame_bytes_against_the_oracle(scene, oracle_big_bunny, big_bunny_inputs):
    """The bench workload itself: bytes of the FRAME kernel's 3840x2160 / 16 rays-per-pixel output against the oracle, at stratified
    pixels over the whole frame and at a dense window over the bunny (x 1760-1920, y 980-1060 at 4K: silhouette, self-shadow)."""
    _, _, smp = big_bunny_inputs
    kw = dict(width=3840, height=2160, nb_ray=16, nb_light_sample=100)
    full, st = scene.render(**kw)
    strat = S.stratified_pixels(3840, 2160, 216, seed=5)                             # 18 x 10 = 180 pixels
    xs, ys = np.meshgrid(np.arange(1760, 1920, 8), np.arange(980, 1060, 8))          # 20 x 10 = 200 pixels on the bunny
    pix = np.concatenate([strat, np.stack([xs.ravel(), ys.ravel()], 1)]).astype(np.uint32)
    o = oracle_big_bunny.render(3840, 2160, smp, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, pixels=pix, nb_ray=16)
    assert (full[pix[:, 1], pix[:, 0]] == o["rgb8"]).all()
    on_bunny = (o["prim0"] >= 0) & (o["prim0"] < 4968)
    assert on_bunny.sum() >= 100, int(on_bunny.sum())


def _multi_scenes(n):
    import rtb200
    import torch
    if torch.cuda.device_count() < n:
        pytest.skip("needs %d GPUs" % n)
    v, c = S.big_bunny()
    smp = S.make_samples()
    scenes = [rtb200.Scene(v, c, device=i) for i in range(n)]
    rtb200.set_samples_multi(scenes, smp)
    return scenes


@pytest.mark.parametrize("n", [2, 4, 8])
def test_render_multi_c_abi_equals_single_gpu(scene, n):
    """rtb_render_multi (one call, one host thread per device, SURVEY 8b n_gpus): bytes == the 1-GPU frame, with every device
    copying its rows to the host frame itself and with the rows gathered over NVLink peer stores into device 0's frame."""
    import rtb200
    scenes = _multi_scenes(n)
    for kw in (dict(width=1920, height=1080, nb_ray=1, nb_light_sample=100), dict(width=3840, height=2160, nb_ray=4, nb_light_sample=25),
               dict(width=333, height=187, distance=60.0, nb_ray=2, nb_light_sample=9, band_rows=8)):
        want, st1 = scene.render(**kw)
        got, st = rtb200.render_multi(scenes, **kw)
        assert (got == want).all(), kw
        assert st["primary_hits"] == st1["primary_hits"] and st["shadow_occluded"] == st1["shadow_occluded"]
        got2, _ = rtb200.render_multi(scenes, gather_peer=True, **kw)
        assert (got2 == want).all(), kw
    # frame-parallel sequence over the devices
    poses = [S.orbit_pose(k) for k in range(24)]
    kw = dict(width=640, height=360, distance=96.0, nb_ray=1, nb_light_sample=16)
    frames, sq = rtb200.render_sequence_multi(scenes, poses, **kw)
    assert sq["frames"] == 24
    for k in range(24):
        eye, look, up = poses[k]
        want, _ = scene.render(eye=eye, look_at=look, up=up, **kw)
        assert (frames[k] == want).all(), k
