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


def env_world() -> Tuple[int, int, int]:
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))


def reduce_accum(accum, dst: int = 0):
    """Sum the (H, W, 4) accumulation tensors of all ranks onto `dst` (in place)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum
