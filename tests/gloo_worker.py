"""world_size-2 gloo worker: each rank renders its strata rows (with the CPU oracle standing in
for the GPU renderer), the linear sums are reduced onto rank 0, rank 0 finalizes and compares
with the unsharded render."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle_binding import OracleWorld  # noqa: E402
from raysnail_b200 import Scene, make_params  # noqa: E402
from raysnail_b200.distributed import env_world, reduce_accum, stratum_rows  # noqa: E402


def main():
    world, rank, _ = env_world()
    dist.init_process_group("gloo")
    scene = Scene.from_sdl(os.path.join(ROOT, "tests", "golden", "sdl", "example.sdl"))
    oracle = OracleWorld(scene)
    w, h, q = 48, 30, 5
    cam = scene.camera(w, h)
    r0, r1 = stratum_rows(rank, world, q)
    n_local = (r1 - r0) * q
    part, _ = oracle.render(cam, make_params(sqrt_spp=q, gamma=False, rows=(r0, r1)), threads=2)
    accum = torch.from_numpy(part.astype(np.float64)[..., :3] * n_local).contiguous()   # back to linear sums
    reduce_accum(accum, 0)
    if rank == 0:
        img = np.sqrt(accum.numpy() / (q * q))
        full, _ = oracle.render(cam, make_params(sqrt_spp=q, gamma=True), threads=2)
        err = np.abs(img - full[..., :3]).max()
        assert err < 2e-6, err
        print("SHARDED_OK", err, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
