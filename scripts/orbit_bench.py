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
    smp = S.make_samples()
    n_workers = int(os.environ.get("ORBIT_WORKERS", "4"))
    mine = list(range(rank, POSES, world))
    frames = np.empty((len(mine), H, W, 3), np.uint8) if "--check" in sys.argv else None

    class Worker:
        """One resident scene + its own streams and double-buffered frames; renders poses mine[w::n_workers].  The camera-grid
        rebuild of a pose synchronises the host a few times; with two workers (host threads, ctypes releases the GIL) one
        scene's kernels fill the gaps the other's rebuild leaves."""
        def __init__(self, w):
            self.sc = rtb200.Scene(v, c, device=local)
            self.sc.set_samples(smp)
            self.idx = list(range(w, len(mine), n_workers))               # positions in `mine`
            self.d_buf = [torch.empty(H * W * 3, dtype=torch.uint8, device=dev) for _ in range(2)]
            self.h_buf = [torch.empty(H * W * 3, dtype=torch.uint8).pin_memory() for _ in range(2)]
            self.render_s, self.copy_s = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
            self.rendered = [torch.cuda.Event() for _ in range(2)]
            self.copied = [torch.cuda.Event() for _ in range(2)]

        def run(self, keep):
            torch.cuda.set_device(local)
            for j, m in enumerate(self.idx):
                b = j & 1
                eye, look, up = S.orbit_pose(mine[m])
                if j >= 2:
                    self.render_s.wait_event(self.copied[b])                  # device buffer b is free again
                    if keep is not None:
                        self.copied[b].synchronize(); keep[self.idx[j - 2]] = self.h_buf[b].numpy().reshape(H, W, 3)
                self.sc.render_device(self.d_buf[b].data_ptr(), self.render_s.cuda_stream, width=W, height=H, eye=eye, look_at=look, up=up)
                self.rendered[b].record(self.render_s)
                self.copy_s.wait_event(self.rendered[b])
                with torch.cuda.stream(self.copy_s):
                    self.h_buf[b].copy_(self.d_buf[b], non_blocking=True)
                self.copied[b].record(self.copy_s)
            self.render_s.synchronize(); self.copy_s.synchronize()
            if keep is not None:
                for j in range(max(0, len(self.idx) - 2), len(self.idx)):
                    keep[self.idx[j]] = self.h_buf[j & 1].numpy().reshape(H, W, 3)

    workers = [Worker(w) for w in range(n_workers)]
    sc = workers[0].sc

    def run(keep):
        import threading
        ts = [threading.Thread(target=wk.run, args=(keep,)) for wk in workers]
        for t in ts: t.start()
        for t in ts: t.join()
        torch.cuda.synchronize(dev)

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
                          "parallelism": "frame-parallel, pose k on rank k mod N, no collective; {} resident scene(s) per rank on their own streams".format(n_workers), "d2h_bytes_per_frame": H * W * 3,
                          "frames_checked_equal_blocking_render": ok}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
