"""world_size-2/3 gloo tests (CPU): the band ownership + equal-stride gather bookkeeping of the multi-GPU path.
The render itself needs a GPU; here each rank fills its compact band buffer with a pattern derived from the GLOBAL
row number, so a wrong row mapping anywhere shows up in the assembled frame."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, width, height, band_rows, q):
    sys.path.insert(0, ROOT)
    import importlib
    import rtb200  # noqa: F401
    D = importlib.import_module("ray_tracer_rust_b200.distributed")
    B = importlib.import_module("ray_tracer_rust_b200.binding")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = D.BandGather(width, height, band_rows, rank, world, "cpu")
        rows = B.owned_row_indices(height, band_rows, rank, world)
        assert len(rows) == g.rows
        pat = ((rows[:, None, None] * 7 + np.arange(width)[None, :, None] * 3 + np.arange(3)[None, None, :]) % 251).astype(np.uint8)
        g.send[: pat.size] = torch.from_numpy(pat.ravel())
        g.collect()
        if rank == 0:
            staged = g.staged.view(world, g.stride).numpy()
            frame = D.assemble_on_host([staged[r] for r in range(world)], width, height, band_rows, world)
            y = np.arange(height)
            want = ((y[:, None, None] * 7 + np.arange(width)[None, :, None] * 3 + np.arange(3)[None, None, :]) % 251).astype(np.uint8)
            q.put(bool((frame == want).all()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,width,height,band_rows,port", [(2, 64, 40, 8, 29611), (3, 50, 77, 8, 29612), (2, 33, 5, 4, 29613)])
def test_band_gather_gloo(built, world, width, height, band_rows, port):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, width, height, band_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
