"""Multi-GPU sharding of one render: the scene is replicated, the STRATUM ROWS
of every pixel are partitioned across ranks, and the per-rank linear radiance sums (float4 per
pixel) are combined with one sum-reduce onto rank 0, which finalizes (divide, gamma).

torch.distributed is plumbing here: NCCL over NVLink on GPUs, gloo in the CPU tests.
"""
from __future__ import annotations

import os
from typing import Tuple


def stratum_rows(rank: int, world_size: int, sqrt_spp: int) -> Tuple[int, int]:
    """Contiguous, balanced split of the strata rows s_j in [0, sqrt_spp) — perfectly load
    balanced (every rank touches every pixel), unlike image tiles.  Ranks beyond sqrt_spp get
    an empty range.  Philox streams are keyed by the absolute (pixel, sample), so the union of
    all ranks' samples is exactly the single-GPU sample set."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    return (sqrt_spp * rank) // world_size, (sqrt_spp * (rank + 1)) // world_size


def pixel_slice(rank: int, world_size: int, n_pixels: int) -> Tuple[int, int]:
    """Pixels [begin, end) that `rank` finalizes in the fused reduce: contiguous, balanced,
    16-pixel aligned (256-byte float4 runs) except for the last rank's tail."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    per = -(-n_pixels // world_size)
    per = (per + 15) // 16 * 16
    return min(n_pixels, rank * per), min(n_pixels, (rank + 1) * per)


class _DevView:
    """Raw device pointer -> torch tensor (zero-copy) through __cuda_array_interface__."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 2, "strides": None}


def device_view(ptr: int, shape, device):
    import torch
    return torch.as_tensor(_DevView(ptr, shape), device=device)


class PeerImage:
    """The per-rank accumulation buffer (and rank 0's output image) as NVLink peer memory.

    Every rank allocates its accumulator through the library (cudaMalloc, so the whole
    allocation is what the IPC handle names), the 64-byte handles are exchanged once with
    all_gather_object, and every rank maps every peer's buffer plus rank 0's image."""

    def __init__(self, renderer, width: int, height: int):
        import torch.distributed as dist
        self.r, self.n_pixels = renderer, width * height
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        nbytes = self.n_pixels * 16
        self.accum = renderer.device_alloc(nbytes)
        self.out = renderer.device_alloc(nbytes) if self.rank == 0 else 0
        import torch
        dev = torch.device("cuda", torch.cuda.current_device())
        self.accum_t = device_view(self.accum, (height, width, 4), dev)  # torch views for zero_() / copies
        self.out_t = device_view(self.out, (height, width, 4), dev) if self.rank == 0 else None
        # arrival counters in rank 0's memory: [0] accumulators complete, [1] image slices written (rs_signal / rs_wait_counter)
        self.counters = renderer.device_alloc(256) if self.rank == 0 else 0
        self.epoch = 0
        mine = {"accum": renderer.ipc_export(self.accum), "out": renderer.ipc_export(self.out) if self.rank == 0 else None,
                "counters": renderer.ipc_export(self.counters) if self.rank == 0 else None}
        handles = [None] * self.world
        dist.all_gather_object(handles, mine)
        self._opened = []
        self.peer_accums = []
        for k, h in enumerate(handles):
            if k == self.rank:
                self.peer_accums.append(self.accum)
            else:
                p = renderer.ipc_open(h["accum"])
                self._opened.append(p)
                self.peer_accums.append(p)
        if self.rank == 0:
            self.out_on_rank0 = self.out
            self.counters_on_rank0 = self.counters
        else:
            self.out_on_rank0 = renderer.ipc_open(handles[0]["out"])
            self.counters_on_rank0 = renderer.ipc_open(handles[0]["counters"])
            self._opened += [self.out_on_rank0, self.counters_on_rank0]

    def reduce_finalize(self, n_samples: int, gamma: bool, stream: int = 0, d_mask=None):
        """Call after this rank's render was enqueued on `stream`; everything here is enqueued on `stream` too
        and nothing blocks the host.  signal(arrived) -> wait(arrived == world) -> fused sum + finalize of this
        rank's pixel slice into rank 0's image -> signal(written) -> wait(written == world), so work queued
        afterwards sees the complete image (rank 0) and may reuse the accumulator (every rank).  The ranks order
        themselves on the devices; there is no host barrier (the two of round 1 are gone)."""
        # the counters stand at world * epoch; epoch moves only after BOTH signals of this rank are queued, so a call that
        # raises before its first signal leaves the count intact (a failure between the two signals cannot be repaired
        # from one process: the peers' waits then end through RS_OPT_WAIT_BUDGET_MS, loudly, instead of spinning for ever)
        expect = self.world * (self.epoch + 1)
        arrived, written = self.counters_on_rank0, self.counters_on_rank0 + 4
        self.r.signal(arrived, stream)
        self.r.wait_counter(arrived, expect, stream)
        b, e = pixel_slice(self.rank, self.world, self.n_pixels)
        self.r.finalize_peers(self.peer_accums, d_mask, self.out_on_rank0, b, e, n_samples, gamma, stream)
        self.r.signal(written, stream)
        self.epoch += 1
        # every rank: nobody reads this rank's accumulator any more (it may be cleared for the next frame), and
        # on rank 0 the image is complete
        self.r.wait_counter(written, expect, stream)

    def close(self):
        # rank 0 owns the counters and the image the other ranks' kernels poll and write: nobody frees anything before
        # every rank has drained its stream and said so
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize()
        if dist.is_initialized() and self.world > 1:
            dist.barrier()
        for p in self._opened:
            self.r.ipc_close(p)
        self._opened = []
        self.r.device_free(self.accum)
        if self.rank == 0 and self.out:
            self.r.device_free(self.out)
        if self.rank == 0 and self.counters:
            self.r.device_free(self.counters)


def env_world() -> Tuple[int, int, int]:
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))


def reduce_accum(accum, dst: int = 0):
    """Sum the (H, W, 4) accumulation tensors of all ranks onto `dst` (in place)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum
