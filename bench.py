#!/usr/bin/env python
"""bench.py -- the path's benchmark (contract in the task statement; extra keys documented in DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config 3|3b|4|5]

--config 3 (default, the headline = BASELINE.json configs[2]): big_bunny 3840x2160, 16 rays/pixel, 100 light samples, reference
camera (distance 288).  3b: the same at the framing-preserving distance 576.  4: 64 flattened big_bunny copies (317 953 primitives),
3840x2160, 1 ray/pixel (configs[3]).  5: 120-pose orbit of big_bunny at 1920x1080, 1 ray/pixel (configs[4]), through
rtb_render_sequence.

A step is one frame (config 5: one orbit of 120 frames).  N > 1 is launched by torchrun, one rank per GPU: the frame's 4-row bands
are interleaved over the ranks (strong scaling: the frame is fixed; config 5: pose k goes to rank k % N, replicas only) and every
rank's shading kernel stores its pixels straight into rank 0's frame over NVLink (PeerFrame); that whole chain is inside the
timed region.  Rank 0 prints ONE JSON line.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time
import zlib

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

NB_LIGHT = 100
BAND_ROWS = 4                 # one tile row per band: finest interleave, per-rank times within 1.2 % at 8 GPUs (8 rows: 6 %)
CONFIGS = {
    "3": dict(width=3840, height=2160, nb_ray=16, distance=288.0, scene="big_bunny",
              workload="big_bunny.obj + ground, 3840x2160, NB_RAY=16, NB_LIGHT_SAMPLE=100, reference camera (distance 288)",
              metric="Mrays/s (primary + shadow rays cast), big_bunny 4K 16 spp"),
    "3b": dict(width=3840, height=2160, nb_ray=16, distance=576.0, scene="big_bunny",
               workload="big_bunny.obj + ground, 3840x2160, NB_RAY=16, NB_LIGHT_SAMPLE=100, framing-preserving camera (distance 576)",
               metric="Mrays/s (primary + shadow rays cast), big_bunny 4K 16 spp, distance 576"),
    "4": dict(width=3840, height=2160, nb_ray=1, distance=288.0, scene="bunnies64",
              workload="64 flattened big_bunny copies (317 953 primitives) + ground, 3840x2160, NB_RAY=1, NB_LIGHT_SAMPLE=100, reference camera",
              metric="Mrays/s (primary + shadow rays cast), 64 bunnies 4K"),
    "5": dict(width=1920, height=1080, nb_ray=1, distance=288.0, scene="big_bunny", poses=120,
              workload="big_bunny.obj + ground, 120-pose orbit, 1920x1080, NB_RAY=1, NB_LIGHT_SAMPLE=100, one step = the 120 frames",
              metric="Mrays/s (primary + shadow rays cast) and frames/s, big_bunny orbit 1080p"),
}


def fp32_flops(st, tri_tests=None, box_tests=None):
    """F_alg = 12 N_box + 46 N_tri + 28 N_primary + 6 N_primary_hit + 35 N_shadow + 15 N_shadow_hit."""
    return (12 * (st["box_tests"] if box_tests is None else box_tests) + 46 * (st["tri_tests"] if tri_tests is None else tri_tests)
            + 28 * st["primary_rays"] + 6 * st["primary_hits"] + 35 * st["shadow_rays"] + 15 * st["shadow_occluded"])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                mx = float(f[1])
                if t0 <= ts <= t1 + 0.2:
                    sm.append(float(f[0]))
                    for n, v in zip(names, f[3:7]):
                        if v.lower().startswith("active"):
                            reasons.add(n)
            except ValueError:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "samples": len(sm), "reasons": sorted(reasons)}


def scene_arrays(cfg):
    import scenes as S
    return S.instanced_bunnies(8) if cfg["scene"] == "bunnies64" else S.big_bunny()


def cpu_baseline_sample(cfg, n_threads=0, target_pixels=None):
    """The oracle in reference-faithful mode (reference tree, no pruning, hardware_concurrency()-1 threads on contiguous slices of a
    shuffled pixel list -- src/main.rs:267-289) on a random pixel subset of the SAME frame.  Config 4: the reference's O(n^2)-per-level
    tree build takes minutes at 317 953 primitives, so the subset is traced by brute force (every primitive per ray), stated in
    `sample`.  Returns (stats dict, seconds, pixels, mode name)."""
    import scenes as S
    from oracle import oracle as O
    verts, colors = scene_arrays(cfg)
    samples = S.make_samples()
    big = cfg["scene"] == "bunnies64"
    sc = O.OracleScene(verts, colors, not big)
    threads = n_threads if n_threads > 0 else max(1, (os.cpu_count() or 2) - 1)
    per_thread = 1500 if cfg["nb_ray"] == 16 else 20000        # ~120 pixels/s/thread at 16 spp x 100 light samples -> ~12 s
    if big:
        per_thread = 200                                        # brute force over 317 953 primitives: ~12 pixels/s/thread -> ~17 s
    n_pix = target_pixels or per_thread * threads
    rng = np.random.default_rng(12345)
    pix = np.stack([rng.integers(0, cfg["width"], n_pix), rng.integers(0, cfg["height"], n_pix)], 1).astype(np.uint32)
    eye, look, up = S.EYE, S.LOOK_AT, S.UP
    r = sc.render(cfg["width"], cfg["height"], samples, eye, look, up, cfg["distance"], S.LIGHT, nb_ray=cfg["nb_ray"],
                  nb_light_sample=NB_LIGHT, mode=O.MODE_BRUTE if big else O.MODE_REFERENCE_TREE, pixels=pix, n_threads=threads)
    return r["stats"], r["stats"]["seconds"], n_pix, ("brute force over all primitives (reference tree not built: O(n^2) per level)" if big
                                                       else "reference tree + unpruned traversal")


def run_reference(args, cfg, rank):
    """--impl reference: the reference's CPU algorithm (C++ restatement: no Rust toolchain here) on the host cores."""
    if rank != 0:
        return
    for _ in range(min(args.warmup, 1)):
        cpu_baseline_sample(cfg, target_pixels=16 if cfg["scene"] == "bunnies64" else 64)
    tot_rays, tot_s, st, n_pix, mode = 0, 0.0, None, 0, ""
    for _ in range(args.steps):
        st, secs, n_pix, mode = cpu_baseline_sample(cfg)
        tot_rays += st["primary_rays"] + st["shadow_rays"]
        tot_s += secs
    value = tot_rays / tot_s / 1e6
    frames = cfg.get("poses", 1)
    frame_rays = frames * cfg["width"] * cfg["height"] * cfg["nb_ray"] * (1 + NB_LIGHT * st["primary_hits"] / max(st["primary_rays"], 1))
    sample = "%d random pixels of the frame per step (%d rays), %s" % (n_pix, st["primary_rays"] + st["shadow_rays"], mode)
    print(json.dumps({
        "impl": "reference", "metric": cfg["metric"], "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": frame_rays / (value * 1e6) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["workload"], "note": "ms_per_step = full-step time extrapolated from the sampled pixels"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": st["threads"], "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def profile_json(name):
    """Figures taken from a committed ncu --set full capture (profiles/<name>): per-launch DRAM / L2 bytes of the hot kernels."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name)))
    except Exception:
        return None


class SharedHostFrame:
    """One host allocation for the whole frame, shared by all ranks of the box (POSIX shared memory) and pinned in each process
    (cudaHostRegister): every rank's rtb_render copies its rows device-to-host straight to their place in it over its own PCIe
    link (RTB_FLAG_FRAME_ROWS), so after a barrier rank 0 holds the complete frame in host memory."""

    def __init__(self, nbytes, rank, world, tag):
        import torch
        from multiprocessing import shared_memory
        self.name = "rtb200_frame_%s" % tag
        self.rank = rank
        self.shm = None
        if rank == 0:
            try:
                shared_memory.SharedMemory(name=self.name).unlink()
            except Exception:
                pass
            self.shm = shared_memory.SharedMemory(name=self.name, create=True, size=nbytes + 64 * world)
            np.ndarray((8 * world,), dtype=np.int64, buffer=self.shm.buf, offset=nbytes)[:] = 0
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=self.name)
        self.array = np.ndarray((nbytes,), dtype=np.uint8, buffer=self.shm.buf)
        self.array[:1] = self.array[:1]
        self.ptr = self.array.ctypes.data
        self.nbytes = nbytes
        # "my rows of step k have landed": one cache line per rank behind the frame, read by everybody (host_barrier)
        self.world = world
        self.landed = np.ndarray((world, 8), dtype=np.int64, buffer=self.shm.buf, offset=nbytes)
        rc = torch.cuda.cudart().cudaHostRegister(self.ptr, nbytes, 0)
        self.registered = int(rc) == 0
        if not self.registered:
            raise SystemExit("cudaHostRegister failed: %r" % (rc,))

    def host_barrier(self, step, timeout_s=60.0):
        """Returns when every rank has finished step `step` (its rows are in the shared frame: rtb_render_end has returned).  A
        host-side rendezvous through the shared allocation itself: the step ends with the complete frame in host memory, and
        nothing about that needs the GPUs (an NCCL barrier costs a kernel launch and a stream synchronisation per step)."""
        self.landed[self.rank, 0] = step
        if self.world == 1:
            return
        t0 = time.perf_counter()
        while int(self.landed[:, 0].min()) < step:
            if time.perf_counter() - t0 > timeout_s:
                raise SystemExit("host_barrier: a rank did not finish step %d" % step)

    def close(self):
        import torch
        self.landed = None
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.ptr)
            self.registered = False
        self.array = None
        try:
            self.shm.close()
            if self.rank == 0:
                self.shm.unlink()
        except Exception:
            pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="3", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, cfg, rank)
        return

    import torch
    import torch.distributed as dist
    import rtb200
    import scenes as S
    D = importlib.import_module("ray_tracer_rust_b200.distributed")

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    W, H, NB_RAY = cfg["width"], cfg["height"], cfg["nb_ray"]
    verts, colors = scene_arrays(cfg)
    samples = S.make_samples()
    scene = rtb200.Scene(verts, colors, device=local_rank)
    scene.set_samples(samples)
    info = scene.info()
    kw = dict(nb_ray=NB_RAY, nb_light_sample=NB_LIGHT, eye=S.EYE, look_at=S.LOOK_AT, up=S.UP, distance=cfg["distance"])
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def allmax(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def allsum_dict(st, keys):
        if world > 1:
            t = torch.tensor([st[k] for k in keys], dtype=torch.int64, device=dev)
            dist.all_reduce(t)
            return dict(zip(keys, [int(x) for x in t.tolist()]))
        return {k: int(st[k]) for k in keys}

    keys = ["primary_rays", "primary_hits", "shadow_rays", "shadow_occluded", "box_tests", "tri_tests"]
    extra = {}
    sampler = None

    if "poses" in cfg:
        # ---------------- config 5: the orbit through rtb_render_sequence; rank r renders poses r, r + N, ... (replicas, no exchange)
        poses_all = [S.orbit_pose(k, cfg["poses"]) for k in range(cfg["poses"])]
        mine = poses_all[rank::world]
        seq_kw = dict(width=W, height=H, nb_ray=NB_RAY, nb_light_sample=NB_LIGHT, distance=cfg["distance"])
        frames = torch.empty((len(mine), H, W, 3), dtype=torch.uint8).pin_memory().numpy()

        def step():
            return scene.render_sequence(mine, frames=frames, **seq_kw)[1]

        for _ in range(max(args.warmup, 3)):
            step()
        barrier()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        time.sleep(0.3 if rank == 0 else 0.0)
        barrier()
        t_wall0 = time.time()
        dev_ms, wall = 0.0, 0.0
        for _ in range(args.steps):
            flush_buf.zero_()
            barrier()
            t0 = time.perf_counter()
            st_seq = step()
            wall += time.perf_counter() - t0
            dev_ms += st_seq["device_ms"]
        barrier()
        t_wall1 = time.time()
        ms = allmax(dev_ms)
        e2e_s = allmax(wall)
        clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
        # ray counts: one counters render per pose of this rank (untimed)
        st = {k: 0 for k in keys}
        for (eye, look, up) in mine:
            _, s1 = scene.render(width=W, height=H, nb_ray=NB_RAY, nb_light_sample=NB_LIGHT, eye=eye, look_at=look, up=up,
                                 distance=cfg["distance"], counters=True)
            for k in keys:
                st[k] += s1[k]
        st = allsum_dict(st, keys)
        # check: frame 0 of this rank's share against a blocking render
        w0, _ = scene.render(width=W, height=H, nb_ray=NB_RAY, nb_light_sample=NB_LIGHT, eye=mine[0][0], look_at=mine[0][1], up=mine[0][2],
                             distance=cfg["distance"])
        extra["sequence_frame_matches_blocking_render"] = bool((frames[0] == w0).all())
        extra["frames_per_s"] = cfg["poses"] / (e2e_s / args.steps)
        extra["frames_per_s_device"] = cfg["poses"] / (ms / args.steps * 1e-3)
        extra["cam_grid_redo_last_step"] = st_seq["cam_grid_redo"]
        h2d, d2h = 64 * len(poses_all), len(poses_all) * H * W * 3
        e2e_path = "rtb_render_sequence: camera poses in, RGB8 frames in pinned host memory out (every frame copied device-to-host)"
        launches_per_step = 0
        kernel_ms, per_kernel, queue, prim_tri, s1 = None, None, None, None, None
        parallelism = "pose k on rank k %% %d (frame-parallel replicas, no exchange)" % world
        use_peer = False
    else:
        # ---------------- configs 3 / 3b / 4: one frame per step, row bands over the ranks
        # N > 1: every rank's shading kernel stores its pixels straight into rank 0's frame over NVLink (CUDA IPC mapping,
        # ray_tracer_rust_b200.distributed.PeerFrame); RTB200_GATHER=nccl selects the send-buffer + NCCL gather + placement form
        use_peer = world > 1 and os.environ.get("RTB200_GATHER", "peer") != "nccl"
        gather = (D.PeerFrame if use_peer else D.BandGather)(W, H, BAND_ROWS, rank, world, dev)

        def step():
            return D.render_distributed(scene, gather, **kw)

        for _ in range(max(args.warmup, 3)):
            step()
        barrier()
        # ---- device-resident timing: K frames, each bracketed by events on the launching stream, L2 flushed between
        sampler = ClockSampler(local_rank) if rank == 0 else None
        time.sleep(0.3 if rank == 0 else 0.0)
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        barrier()
        t_wall0 = time.time()
        for i in range(args.steps):
            flush_buf.zero_()
            if world > 1:
                dist.barrier()
            starts[i].record(stream)
            frame_dev = step()
            ends[i].record(stream)
        barrier()
        t_wall1 = time.time()
        ms = allmax(sum(s.elapsed_time(e) for s, e in zip(starts, ends)))
        clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
        # the gathered frame (rank 0's device memory, written by all ranks over NVLink) against rank 0 rendering the frame alone
        if rank == 0:
            got = frame_dev.cpu().numpy().reshape(H, W, 3)
            alone, _ = scene.render(width=W, height=H, **kw)
            extra["frame_matches_single_gpu"] = bool((got == alone).all())
            extra["frame_crc32"] = "%08x" % (zlib.crc32(got.tobytes()) & 0xffffffff)
        if world > 1:
            dist.barrier()

        # ---- ray counts + traversal counters of this rank's bands (instrumented kernels, untimed), summed over ranks
        _, st = scene.render(width=W, height=H, band_rows=BAND_ROWS, rank=rank, world=world, counters=True, **kw)
        st = allsum_dict(st, keys)
        _, stp = scene.render(width=W, height=H, band_rows=BAND_ROWS, rank=rank, world=world, counters=True, **dict(kw, nb_light_sample=0))
        prim_tri = allsum_dict(stp, keys)          # nb_light_sample = 0: only k_primary runs -> its own box / triangle tests
        # kernel-only times of this rank's chain (events inside the library; chunks in sequence on one stream)
        kms, per_kernel = [], {"k_primary": [], "k_shadow_grid": [], "k_shade": []}
        for _ in range(3):
            flush_buf.zero_()
            torch.cuda.synchronize(dev)
            _, s1 = scene.render(width=W, height=H, band_rows=BAND_ROWS, rank=rank, world=world, profile=True, **kw)
            kms.append(s1["kernel_ms"])
            per_kernel["k_primary"].append(s1["primary_ms"])
            per_kernel["k_shadow_grid"].append(s1["shadow_ms"])
            per_kernel["k_shade"].append(s1["shade_ms"])
        kernel_ms = float(np.mean(kms))
        per_kernel = {k: float(np.mean(v)) for k, v in per_kernel.items()}
        queue = {"queued_hit_points": s1["queued_hits"], "path": "light grids" if s1["grid_cells"] else "candidate lists / BVH walks",
                 "grid_cells": s1["grid_cells"], "grid_entries": s1["grid_entries"], "grid_big_primitives": s1["grid_big"],
                 "grid_build_ms_per_sample_table": s1["grid_build_ms"], "list_pairs": s1["queued_groups"]}

        # ---- end to end through the public C ABI with HOST buffers, every step: asynchronous H2D of the 16 MB sample table from
        #      pinned memory (the reference draws a fresh table per render(), src/main.rs:260-265; a new table makes the library
        #      rebuild its light grids), the frame, and D2H of every rank's rows straight to their place in ONE host frame that
        #      all ranks share (SharedHostFrame): after the barrier rank 0 holds the complete frame in host memory.
        tag = os.environ.get("MASTER_PORT", "0") + "_" + str(os.getppid() if world > 1 else os.getpid())
        shared = SharedHostFrame(H * W * 3, rank, world, tag)
        out_np = shared.array.reshape(H, W, 3)
        # two pinned copies of the table, used in turn: the copy a frame in flight reads is not the one handed over for the next
        h_smp = [torch.from_numpy(samples).pin_memory() for _ in range(2)]
        smp_np = [h.numpy() for h in h_smp]
        rkw = dict(width=W, height=H, band_rows=BAND_ROWS, rank=rank, world=world, frame_rows=(world > 1), **kw)
        n_e2e = [0]

        def e2e_step():
            # the step's table was handed over during the previous step (or just before the loop): its H2D copy, light tables and
            # light grids ran next to that step's kernels; this step does the same for the next one
            k = n_e2e[0]; n_e2e[0] += 1
            scene.render_begin(out_np, **rkw)
            scene.set_samples_async(smp_np[(k + 1) & 1])
            se = scene.render_end()[1]
            shared.host_barrier(k + 1)               # "every rank's rows have landed": the step ends with the full frame on the host
            return se

        scene.set_samples_async(smp_np[0])
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            se = e2e_step()
        barrier()
        e2e_s = allmax(time.perf_counter() - t0)
        if rank == 0:
            out_np[:] = 0
        barrier()
        se = e2e_step()
        barrier()
        if rank == 0:
            alone, _ = scene.render(width=W, height=H, **kw)
            extra["e2e_host_frame_matches_single_gpu"] = bool((out_np == alone).all())
            extra["e2e_grid_rebuilds_per_step"] = se["grid_rebuilds"]
            extra["e2e_grid_redo_last_step"] = se["grid_redo"]
        barrier()
        shared.close()
        h2d, d2h = world * int(samples.nbytes + 64), H * W * 3
        e2e_path = ("per rank and step: rtb_render_begin (RTB_FLAG_FRAME_ROWS) / rtb_scene_set_samples_async (the NEXT step's pinned table: "
                    "H2D + light-grid rebuild next to this step's kernels) / rtb_render_end; every rank's rows copied device-to-host into one "
                    "shared pinned host frame, host-side rendezvous per step; light grids rebuilt every step (new table)")
        launches_per_step = s1["launches"] * world + (1 if (world > 1 and not use_peer) else 0)
        parallelism = ("row bands b %% N over %d GPU(s), scene replicated, " % world) + (
            "k_shade stores into rank 0's frame over NVLink peer memory (one kernel renders and gathers)" if use_peer
            else ("no exchange" if world == 1 else "NCCL gather to rank 0"))

    if rank == 0:
        rays = st["primary_rays"] + st["shadow_rays"]
        ms_per_step = ms / args.steps
        value = rays / (ms_per_step * 1e-3) / 1e6
        e2e_value = rays / (e2e_s / args.steps) / 1e6
        props = torch.cuda.get_device_properties(dev)
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            peaks = {}
        measured = rtb200.measure_peaks(local_rank)                   # k_fp32_peak / k_l2_peak on this device, now
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0
        theoretical = props.multi_processor_count * 128 * 2 * sm_mhz * 1e6 / 1e12
        peak_tflops = measured["fp32_tflops"] if measured["fp32_tflops"] > 0 else theoretical
        flops = fp32_flops(st)
        # F_alg / (t_frame * P_fp32) on the frame time; at N ranks every GPU does 1/N of the flops in that time
        achieved = flops / world / (ms_per_step * 1e-3) / 1e12
        roofline = {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                    "traffic": profile_json("r2_traffic.json"),
                    "kernel": "wavefront chain k_primary -> k_shadow_grid -> k_shade per chunk of 16 Mi rays; frac = F_alg / (ms_per_step x peak)",
                    "algorithmic_flops_per_launch": flops / world,
                    "peak_source": "k_fp32_peak measured in this run (%.1f TFLOP/s; theoretical %d SMs x 128 lanes x 2 x %.0f MHz = %.1f)"
                                   % (measured["fp32_tflops"], props.multi_processor_count, sm_mhz, theoretical),
                    "l2_peak_gbs": measured["l2_read_gbs"], "hbm_read_gbs": measured["hbm_read_gbs"],
                    "hbm_gbs_measured_peaks_json": peaks.get("hbm_gbs")}
        if per_kernel is not None:
            # per-kernel fractions (RTB_FLAG_PROFILE: chunks in sequence on one stream), each with its own share of F_alg:
            # k_primary: ray generation + its triangle / box tests + hit points; k_shadow_grid: the shadow rays' tests and occluder
            # distances; k_shade: the 35 flops of every shadow ray (direction, normalise, |n.l|, accumulate)
            # (the primary-only render has no light grids, so its k_primary plans every hit point's shadow rays over the BVH: those box
            # tests are not part of the frame -- the frame's own counters bound what k_primary can be credited with)
            p_box = min(prim_tri["box_tests"], st["box_tests"]); p_tri = min(prim_tri["tri_tests"], st["tri_tests"])
            f_primary = 28 * st["primary_rays"] + 6 * st["primary_hits"] + 46 * p_tri + 12 * p_box
            f_shadow = 46 * (st["tri_tests"] - p_tri) + 12 * (st["box_tests"] - p_box) + 15 * st["shadow_occluded"]
            f_shade = 35 * st["shadow_rays"]
            pk = {}
            for name, f in (("k_primary", f_primary), ("k_shadow_grid", f_shadow), ("k_shade", f_shade)):
                t = per_kernel[name]
                pk[name] = {"ms": t, "flops": f / world, "tflops": (f / world / (t * 1e-3) / 1e12) if t > 0 else None,
                            "frac": (f / world / (t * 1e-3) / 1e12 / peak_tflops) if t > 0 else None}
            l2 = profile_json("r2_l2_bytes.json")                     # lts__t_bytes per launch from the committed ncu captures
            if l2:
                for name in pk:
                    if name in l2 and pk[name]["ms"]:
                        pk[name]["l2_gbs_ncu_capture"] = l2[name].get("l2_gbs")
                        pk[name]["l2_frac_of_measured_peak"] = (l2[name].get("l2_gbs") or 0) / measured["l2_read_gbs"] if measured["l2_read_gbs"] else None
            roofline["per_kernel"] = pk
            roofline["kernel_ms_serialised"] = kernel_ms
            roofline["shadow_queue"] = queue
        out = {
            "metric": cfg["metric"], "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": cfg["workload"], "primitives": info["n_tris"], "bvh_nodes": info["n_nodes"], "band_rows": BAND_ROWS,
                       "parallelism": parallelism, "l2": "flushed between timed steps (256 MiB write)",
                       "sample_table": "SplitMix64 seed 0x5EED0001, 2e6 pairs", "tie_mode": info["tie_mode"],
                       "light_grids": "`value`: the table is resident, so its grids are in their reused-table form (lists fully ordered on the 2nd "
                                      "frame, cells refined to 0.1 extents on the 3rd: both inside the warm-up); `e2e`: a new table per step, "
                                      "quick build (0.25 extents, two list classes) next to the previous step's kernels"},
            "mrays_per_s_primary_only": st["primary_rays"] / (ms_per_step * 1e-3) / 1e6,
            "rays_per_step": {"primary": st["primary_rays"], "primary_hits": st["primary_hits"], "shadow": st["shadow_rays"],
                              "shadow_occluded": st["shadow_occluded"]},
            "traversal": {"box_tests_per_ray": st["box_tests"] / rays, "tri_tests_per_ray": st["tri_tests"] / rays,
                          "sah_cost": info["sah_cost"], "bvh_depth": info["max_depth"]},
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_s / args.steps * 1e3,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "path": e2e_path},
            "gpu_launches": args.steps * launches_per_step if launches_per_step else None,
            "roofline": roofline,
            "clocks": clocks,
        }
        out.update(extra)
        if out["gpu_launches"] is None:                               # config 5: per frame k_primary + k_shadow_grid + k_shade (+ terms) + the camera-grid build
            out["gpu_launches"] = args.steps * cfg["poses"] * 16
        if not args.no_cpu_baseline and world == 1:
            cst, secs, n_pix, mode = cpu_baseline_sample(cfg)
            cval = (cst["primary_rays"] + cst["shadow_rays"]) / secs / 1e6
            out["cpu_baseline"] = {"value": cval, "unit": "Mrays/s", "cores": cst["threads"], "kind": "port",
                                   "sample": "%d random pixels of the same frame (%d rays, %.1f s), C++ restatement of the "
                                             "Rust reference: %s" % (n_pix, cst["primary_rays"] + cst["shadow_rays"], secs, mode)}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        if "poses" not in cfg and hasattr(gather, "close"):
            gather.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
