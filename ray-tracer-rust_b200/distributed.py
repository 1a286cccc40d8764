"""Multi-GPU frame assembly: one process per GPU, scene replicated, interleaved row bands, one gather to rank 0.

Replaces the thread fan-out of render() (src/main.rs:269-303): where the reference gives each CPU thread a slice
of a shuffled pixel list and serialises put_pixel behind a Mutex, here rank r renders every band b with
b % world == r straight into its NCCL send buffer (rtb_render_device), the buffers are gathered to rank 0 over
NVLink, and rank 0 runs the de-interleave kernel (rtb_deinterleave_bands).  Pixels are independent, so there is no
other exchange.  torch / torch.distributed are used for device memory, streams and the collective only.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import binding


def padded_stride(width, height, band_rows, world):
    """Bytes per rank in the gather: the largest band set, rounded up to 16 B (uint4 copies, equal-size gather)."""
    rows = max(binding.owned_rows(height, band_rows, r, world) for r in range(world))
    return (rows * width * 3 + 15) // 16 * 16


class BandGather:
    """Owns the send buffer (all ranks), the staging buffer and the final frame (rank 0)."""

    def __init__(self, width, height, band_rows, rank, world, device, group=None):
        self.width, self.height, self.band_rows, self.rank, self.world = width, height, band_rows, rank, world
        self.device = torch.device(device)
        self.group = group
        self.stride = padded_stride(width, height, band_rows, world)
        self.rows = binding.owned_rows(height, band_rows, rank, world)
        self.send = torch.zeros(self.stride, dtype=torch.uint8, device=self.device)
        self.staged = None
        self.frame = None
        if rank == 0:
            self.staged = torch.zeros(world * self.stride, dtype=torch.uint8, device=self.device)
            self.frame = torch.zeros(height * width * 3, dtype=torch.uint8, device=self.device)

    def collect(self):
        """The collective alone: per-rank compact buffers -> rank 0's staging buffer (world x stride)."""
        if self.rank == 0:
            parts = list(self.staged.view(self.world, self.stride).unbind(0))
            dist.gather(self.send, parts, dst=0, group=self.group)
        else:
            dist.gather(self.send, None, dst=0, group=self.group)

    def gather(self):
        """Collective + placement: after it returns (stream-ordered) rank 0's self.frame holds the row-major RGB8
        frame.  world == 1 needs neither: the send buffer already is the frame."""
        if self.world == 1:
            self.frame = self.send[: self.height * self.width * 3]
            return self.frame
        self.collect()
        if self.rank != 0:
            return None
        if not self.staged.is_cuda:
            raise binding.RtbError("de-interleave runs on the GPU only (no CPU path)")
        binding.deinterleave_bands(self.device.index or 0, self.staged.data_ptr(), self.stride, self.frame.data_ptr(),
                                   self.width, self.height, self.band_rows, self.world,
                                   torch.cuda.current_stream(self.device).cuda_stream)
        return self.frame


class _DevicePtr:
    """A raw device allocation as an object torch.as_tensor can wrap (__cuda_array_interface__)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerFrame:
    """Render + gather as ONE kernel: rank 0 owns the frame, every other rank maps it through CUDA IPC (NVLink peer
    access) and its shading kernel stores each pixel straight into rank 0's frame at the pixel's own position
    (RTB_FLAG_FRAME_ROWS).  No send buffers, no collective moving data, no placement pass; the only cross-rank step per
    frame is a one-element all_reduce that orders "every rank's kernel has finished" before rank 0 reads the frame.
    TWO frames alternate: while rank 0 still reads (copies, encodes) frame N from one buffer, the peers' kernels of frame
    N + 1 store into the other one; the all_reduce of frame N + 1 -- which rank 0 enqueues after its reads of frame N - 1 on
    the same stream -- is what orders "rank 0 has finished with buffer (N + 1) % 2" before anybody's frame N + 2 writes it.
    Same interface as BandGather (send / gather() / frame)."""

    def __init__(self, width, height, band_rows, rank, world, device, group=None):
        self.width, self.height, self.band_rows, self.rank, self.world = width, height, band_rows, rank, world
        self.device = torch.device(device)
        self.group = group
        self.nbytes = height * width * 3
        self.index = self.device.index or 0
        self.owner = rank == 0
        self.ptrs, self.frames = [0, 0], [None, None]
        for k in range(2):
            handle = torch.zeros(64, dtype=torch.uint8, device=self.device)
            if self.owner:
                self.ptrs[k], h = binding.peer_alloc(self.index, self.nbytes)
                handle.copy_(torch.frombuffer(bytearray(h), dtype=torch.uint8))
            if world > 1:
                dist.broadcast(handle, src=0, group=group)
            if not self.owner:
                self.ptrs[k] = binding.peer_open(self.index, bytes(handle.cpu().numpy().tobytes()))
            if self.owner:
                self.frames[k] = torch.as_tensor(_DevicePtr(self.ptrs[k], self.nbytes), device=self.device)
        self.cur = 0
        self.flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.frame_rows = True

    @property
    def ptr(self):
        """The buffer the next frame is rendered into."""
        return self.ptrs[self.cur]

    @property
    def frame(self):
        return self.frames[self.cur]

    def gather(self):
        """Stream-ordered: after it, the returned tensor (rank 0) holds the row-major RGB8 frame of every rank's rows; the next
        frame goes to the other buffer."""
        if self.world > 1:
            dist.all_reduce(self.flag, group=self.group)
        done = self.frames[self.cur]
        self.cur ^= 1
        return done

    def close(self):
        if self.ptrs[0]:
            torch.cuda.synchronize(self.device)
            if self.world > 1:
                dist.barrier(group=self.group)           # nobody unmaps / frees while a peer may still write
            self.frames = [None, None]
            for k in range(2):
                binding.peer_close(self.index, self.ptrs[k], self.owner)
            self.ptrs = [0, 0]


def render_distributed(scene, gather, **params):
    """One frame on this rank's bands + the gather.  Asynchronous on the current torch stream of the scene's device."""
    stream = torch.cuda.current_stream(gather.device).cuda_stream
    if getattr(gather, "frame_rows", False):             # PeerFrame: pixels go straight into rank 0's frame
        scene.render_device(gather.ptr, stream, width=gather.width, height=gather.height, band_rows=gather.band_rows,
                            rank=gather.rank, world=gather.world, frame_rows=True, **params)
    else:
        scene.render_device(gather.send.data_ptr(), stream, width=gather.width, height=gather.height,
                            band_rows=gather.band_rows, rank=gather.rank, world=gather.world, **params)
    return gather.gather()


def assemble_on_host(per_rank_rows, width, height, band_rows, world):
    """Bookkeeping mirror of the de-interleave for host-side tests: list of per-rank compact byte buffers (possibly
    padded) -> (height, width, 3) frame."""
    frame = np.zeros((height, width, 3), np.uint8)
    for r in range(world):
        idx = binding.owned_row_indices(height, band_rows, r, world)
        buf = np.asarray(per_rank_rows[r], np.uint8).ravel()[: len(idx) * width * 3]
        frame[idx] = buf.reshape(len(idx), width, 3)
    return frame
