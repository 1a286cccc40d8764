#!/usr/bin/env python
"""bench.py -- headline benchmark: big_bunny at 3840x2160, 16 rays/pixel, 100 light samples (BASELINE.json configs[2]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step is one frame.  N > 1 is launched by torchrun (one rank per GPU): the frame's row bands are interleaved over the
ranks (strong scaling: the frame is fixed), gathered to rank 0 with NCCL and de-interleaved there; that whole chain is
inside the timed region.  Rank 0 prints ONE JSON line (contract in the task statement; extra keys documented in
DESIGN.md section "Measurement").
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

WIDTH, HEIGHT, NB_RAY, NB_LIGHT = 3840, 2160, 16, 100
WORKLOAD = "big_bunny.obj + ground, 3840x2160, NB_RAY=16, NB_LIGHT_SAMPLE=100, reference camera (distance 288)"
METRIC = "Mrays/s (primary + shadow rays cast), big_bunny 4K 16 spp"
BAND_ROWS = 4                 # one tile row per band: finest interleave, per-rank times within 1.2 % at 8 GPUs (8 rows: 6 %)


def fp32_flops(st):
    """F_alg = 12 N_box + 46 N_tri + 28 N_primary + 6 N_primary_hit + 35 N_shadow + 15 N_shadow_hit."""
    return (12 * st["box_tests"] + 46 * st["tri_tests"] + 28 * st["primary_rays"] + 6 * st["primary_hits"]
            + 35 * st["shadow_rays"] + 15 * st["shadow_occluded"])


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


def cpu_baseline_sample(n_threads=0, target_pixels=None):
    """The oracle in reference-faithful mode (reference tree, no pruning, hardware_concurrency()-1 threads on
    contiguous slices of a shuffled pixel list -- src/main.rs:267-289) on a random pixel subset of the SAME 4K/16spp
    frame.  Returns (stats dict, seconds, pixels)."""
    import scenes as S
    from oracle import oracle as O
    verts, colors = S.big_bunny()
    samples = S.make_samples()
    sc = O.OracleScene(verts, colors, True)
    threads = n_threads if n_threads > 0 else max(1, (os.cpu_count() or 2) - 1)
    n_pix = target_pixels or 1500 * threads     # ~120 pixels/s/thread at 16 spp x 100 light samples -> ~12 s
    rng = np.random.default_rng(12345)
    pix = np.stack([rng.integers(0, WIDTH, n_pix), rng.integers(0, HEIGHT, n_pix)], 1).astype(np.uint32)
    r = sc.render(WIDTH, HEIGHT, samples, S.EYE, S.LOOK_AT, S.UP, S.DISTANCE, S.LIGHT, nb_ray=NB_RAY,
                  nb_light_sample=NB_LIGHT, mode=O.MODE_REFERENCE_TREE, pixels=pix, n_threads=threads)
    return r["stats"], r["stats"]["seconds"], n_pix


def run_reference(args, rank):
    """--impl reference: the reference's CPU algorithm (C++ restatement: no Rust toolchain here) on the host cores."""
    if rank != 0:
        return
    for _ in range(min(args.warmup, 1)):
        cpu_baseline_sample(target_pixels=64)
    tot_rays, tot_s, st, n_pix = 0, 0.0, None, 0
    for _ in range(args.steps):
        st, secs, n_pix = cpu_baseline_sample()
        tot_rays += st["primary_rays"] + st["shadow_rays"]
        tot_s += secs
    value = tot_rays / tot_s / 1e6
    # one full frame at this rate
    frame_rays = WIDTH * HEIGHT * NB_RAY * (1 + NB_LIGHT * st["primary_hits"] / st["primary_rays"])
    sample = "%d random pixels of the 4K/16spp frame per step (%d rays), reference tree + unpruned traversal" % (
        n_pix, st["primary_rays"] + st["shadow_rays"])
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": frame_rays / (value * 1e6) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "ms_per_step = full-frame time extrapolated from the sampled pixels"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": st["threads"], "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def traffic_of_profile():
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the chain's largest kernel from the committed ncu --set full
    capture (profiles/r1d_traffic.json names the kernel, the launch and the report); None when absent."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r1d_traffic.json")))
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
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

    verts, colors = S.big_bunny()
    samples = S.make_samples()
    scene = rtb200.Scene(verts, colors, device=local_rank)
    scene.set_samples(samples)
    info = scene.info()
    kw = dict(nb_ray=NB_RAY, nb_light_sample=NB_LIGHT, eye=S.EYE, look_at=S.LOOK_AT, up=S.UP, distance=S.DISTANCE)
    # N > 1: every rank's shading kernel stores its pixels straight into rank 0's frame over NVLink (CUDA IPC mapping,
    # ray_tracer_rust_b200.distributed.PeerFrame); RTB200_GATHER=nccl selects the send-buffer + NCCL gather + placement form
    use_peer = world > 1 and os.environ.get("RTB200_GATHER", "peer") != "nccl"
    gather = (D.PeerFrame if use_peer else D.BandGather)(WIDTH, HEIGHT, BAND_ROWS, rank, world, dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step():
        D.render_distributed(scene, gather, **kw)

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
        step()
        ends[i].record(stream)
    barrier()
    t_wall1 = time.time()
    ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None

    # ---- ray counts + traversal counters of this rank's bands (instrumented kernel, untimed), summed over ranks
    _, st = scene.render(width=WIDTH, height=HEIGHT, band_rows=BAND_ROWS, rank=rank, world=world, counters=True, **kw)
    keys = ["primary_rays", "primary_hits", "shadow_rays", "shadow_occluded", "box_tests", "tri_tests"]
    if world > 1:
        t = torch.tensor([st[k] for k in keys], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        st = dict(zip(keys, [int(x) for x in t.tolist()]))
    # kernel-only time of this rank's trace kernel (events inside the library, stream of the launch)
    kms, per_kernel = [], {"k_primary": [], "k_shadow_grid": [], "k_shade": []}
    for _ in range(3):
        flush_buf.zero_()
        torch.cuda.synchronize(dev)
        _, s1 = scene.render(width=WIDTH, height=HEIGHT, band_rows=BAND_ROWS, rank=rank, world=world, profile=True, **kw)
        kms.append(s1["kernel_ms"])
        per_kernel["k_primary"].append(s1["primary_ms"])
        per_kernel["k_shadow_grid"].append(s1["shadow_ms"])
        per_kernel["k_shade"].append(s1["shade_ms"])
    kernel_ms = float(np.mean(kms))
    per_kernel = {k: float(np.mean(v)) for k, v in per_kernel.items()}
    queue = {"queued_hit_points": s1["queued_hits"], "path": "light grids" if s1["grid_cells"] else "candidate lists / BVH walks",
             "grid_cells": s1["grid_cells"], "grid_entries": s1["grid_entries"], "grid_big_primitives": s1["grid_big"],
             "grid_build_ms_per_sample_table": s1["grid_build_ms"], "list_pairs": s1["queued_groups"]}

    # ---- end to end through the public C ABI with HOST buffers: per step H2D of the sample table (the reference draws
    #      a fresh table per render(), src/main.rs:260-265) + params, render, D2H of this rank's RGB8 rows; N>1: the rows
    #      are then gathered on the host side by rank 0 (gloo-free: via NCCL on a device copy would hide the D2H)
    rows = rtb200.owned_rows(HEIGHT, BAND_ROWS, rank, world)
    h_out = torch.empty(rows * WIDTH * 3, dtype=torch.uint8).pin_memory()
    h_smp = torch.from_numpy(samples).pin_memory()
    out_np, smp_np = h_out.numpy().reshape(rows, WIDTH, 3), h_smp.numpy()

    def e2e_step():
        scene.set_samples(smp_np)
        scene.render(out=out_np, width=WIDTH, height=HEIGHT, band_rows=BAND_ROWS, rank=rank, world=world, **kw)

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())

    if rank == 0:
        rays = st["primary_rays"] + st["shadow_rays"]
        ms_per_step = ms / args.steps
        value = rays / (ms_per_step * 1e-3) / 1e6
        e2e_value = rays / (e2e_s / args.steps) / 1e6
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            peaks = {}
        props = torch.cuda.get_device_properties(dev)
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0
        peak_tflops = props.multi_processor_count * 128 * 2 * sm_mhz * 1e6 / 1e12
        flops = fp32_flops(st)
        # per-launch: the dominant kernel is k_trace_frame, one launch per frame per rank; at N ranks a launch does 1/N
        achieved = flops / world / (kernel_ms * 1e-3) / 1e12
        out = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "primitives": info["n_tris"], "bvh_nodes": info["n_nodes"], "band_rows": BAND_ROWS,
                       "parallelism": ("row bands b %% N over %d GPU(s), scene replicated, " % world) + ("k_shade stores into rank 0's frame over NVLink peer memory (one kernel renders and gathers)" if use_peer else ("no exchange" if world == 1 else "NCCL gather to rank 0")),
                       "l2": "flushed between timed frames (256 MiB write)", "sample_table": "SplitMix64 seed 0x5EED0001, 2e6 pairs"},
            "mrays_per_s_primary_only": st["primary_rays"] / (ms_per_step * 1e-3) / 1e6,
            "rays_per_frame": {"primary": st["primary_rays"], "primary_hits": st["primary_hits"], "shadow": st["shadow_rays"],
                               "shadow_occluded": st["shadow_occluded"]},
            "traversal": {"box_tests_per_ray": st["box_tests"] / rays, "tri_tests_per_ray": st["tri_tests"] / rays,
                          "sah_cost": info["sah_cost"], "bvh_depth": info["max_depth"]},
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": e2e_s / args.steps * 1e3,
                    "h2d_bytes_per_step": int(samples.nbytes + 64), "d2h_bytes_per_step": int(rows * WIDTH * 3),
                    "path": "rtb_scene_set_samples + rtb_render with pinned host buffers"},
            "gpu_launches": args.steps * (s1["launches"] * world + (1 if (world > 1 and not use_peer) else 0)),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                         "traffic": traffic_of_profile(), "kernel": "wavefront chain k_primary -> k_shadow_grid -> k_shade per chunk of 16 Mi rays (per_kernel_ms: chunks in sequence, RTB_FLAG_PROFILE)",
                         "kernel_ms": kernel_ms, "per_kernel_ms": per_kernel, "shadow_queue": queue, "algorithmic_flops_per_launch": flops / world,
                         "peak_source": "SMs x 128 FP32 lanes x 2 x SM clock under load (%d SMs, %.0f MHz); MEASURED_PEAKS.json has no FP32 entry"
                                        % (props.multi_processor_count, sm_mhz)},
            "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            cst, secs, n_pix = cpu_baseline_sample()
            cval = (cst["primary_rays"] + cst["shadow_rays"]) / secs / 1e6
            out["cpu_baseline"] = {"value": cval, "unit": "Mrays/s", "cores": cst["threads"], "kind": "port",
                                   "sample": "%d random pixels of the same 4K/16spp frame (%d rays, %.1f s), C++ restatement of the "
                                             "Rust reference: reference tree + unpruned traversal" % (n_pix, cst["primary_rays"] + cst["shadow_rays"], secs)}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        if hasattr(gather, "close"):
            gather.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
