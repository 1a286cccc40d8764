"""BASELINE config 5: frame-sequence orbit of big_bunny (120 camera poses, 1920x1080, NB_RAY=1, 100 light samples),
sustained frames/s with every frame delivered to pinned HOST memory (the reference's render -> image tail without the
PNG encode, src/main.rs:291-315).  Camera-only updates: the scene, BVH, sample table and light grids stay resident,
only the camera grid is rebuilt per pose.

Pipeline per rank: frame k renders on the render stream into device buffer k % 2 (rtb_render_device, asynchronous);
a copy stream moves it to pinned host buffer k % 2 while frame k + 1 renders.  N GPUs: frame-parallel, pose k goes
to rank k % N, no collective ("replicas only": poses are independent).  Time = max over ranks, barrier on both sides.

usage: python scripts/orbit_bench.py            (1 GPU)
       python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/orbit_bench.py
Prints one JSON line on rank 0.  With --check the frames of poses 0 and 60 are compared with blocking rtb_render calls.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist

import rtb200
import scenes as S

W, H, POSES = 1920, 1080, 120


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    v, c = S.big_bunny()
    sc = rtb200.Scene(v, c, device=local)
    sc.set_samples(S.make_samples())
    mine = list(range(rank, POSES, world))
    d_buf = [torch.empty(H * W * 3, dtype=torch.uint8, device=dev) for _ in range(2)]
    h_buf = [torch.empty(H * W * 3, dtype=torch.uint8).pin_memory() for _ in range(2)]
    frames = np.empty((len(mine), H, W, 3), np.uint8) if "--check" in sys.argv else None
    render_s, copy_s = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    rendered = [torch.cuda.Event() for _ in range(2)]
    copied = [torch.cuda.Event() for _ in range(2)]

    def run(keep):
        for j, k in enumerate(mine):
            b = j & 1
            eye, look, up = S.orbit_pose(k)
            if j >= 2:
                render_s.wait_event(copied[b])                       # device buffer b is free again
                if keep is not None:
                    copied[b].synchronize(); keep[j - 2] = h_buf[b].numpy().reshape(H, W, 3)
            sc.render_device(d_buf[b].data_ptr(), render_s.cuda_stream, width=W, height=H, eye=eye, look_at=look, up=up)
            rendered[b].record(render_s)
            copy_s.wait_event(rendered[b])
            with torch.cuda.stream(copy_s):
                h_buf[b].copy_(d_buf[b], non_blocking=True)
            copied[b].record(copy_s)
        torch.cuda.synchronize(dev)
        if keep is not None:
            for j in range(max(0, len(mine) - 2), len(mine)):
                keep[j] = h_buf[j & 1].numpy().reshape(H, W, 3)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    run(None)                                                        # warm-up: all buffers and grids sized
    barrier()
    t0 = time.perf_counter()
    run(frames)
    barrier()
    secs = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([secs], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
    ok = None
    if frames is not None:
        ok = True
        for k in (0, 60):
            if k in mine:
                eye, look, up = S.orbit_pose(k)
                want, _ = sc.render(width=W, height=H, eye=eye, look_at=look, up=up)
                ok = ok and bool((frames[mine.index(k)] == want).all())
    if rank == 0:
        print(json.dumps({"metric": "frames/s, 120-pose orbit of big_bunny, 1920x1080, NB_RAY=1, 100 light samples, frames delivered to pinned host memory",
                          "value": POSES / secs, "unit": "frames/s", "n_gpus": world, "ms_per_frame": secs / POSES * 1e3,
                          "parallelism": "frame-parallel, pose k on rank k % N, no collective", "d2h_bytes_per_frame": H * W * 3,
                          "frames_checked_equal_blocking_render": ok}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
