"""Multi-GPU correctness check (run under torchrun on a box with N GPUs; not a pytest file):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/multi_gpu_check.py

Every rank renders its strata rows of the book-1 final scene; the image combined by the fused
peer-memory reduce + finalize kernel (rs_finalize_peers) must equal (a) the image combined by
ncclReduce + rs_finalize and (b) the image rank 0 renders alone, to f32 summation-order noise.
Prints one JSON line on rank 0 and exits non-zero on a mismatch.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from raysnail_b200 import Renderer, Scene, make_params  # noqa: E402
from raysnail_b200.distributed import PeerImage, reduce_accum, stratum_rows  # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    w, h, q = 480, 270, 8
    scene = Scene.builtin("rtow_final", 7)
    r = Renderer(local)
    r.upload(scene)
    cam = scene.camera(w, h)
    rows = stratum_rows(rank, world, q)
    # (1) fused peer path
    peer = PeerImage(r, w, h)
    peer.accum_t.zero_()
    r.render_accumulate(cam, make_params(sqrt_spp=q, rows=rows), None, peer.accum, 0)
    peer.reduce_finalize(q * q, True, 0)
    # (2) NCCL baseline
    acc = torch.zeros((h, w, 4), dtype=torch.float32, device=dev)
    r.render_accumulate(cam, make_params(sqrt_spp=q, rows=rows), None, acc.data_ptr(), 0)
    torch.cuda.synchronize()
    reduce_accum(acc, 0)
    ok = True
    line = {"world_size": world}
    if rank == 0:
        out_nccl = torch.empty_like(acc)
        r.finalize(acc.data_ptr(), None, out_nccl.data_ptr(), w, h, q * q, True, 0)
        # (3) single GPU
        solo = torch.zeros_like(acc)
        r.render_accumulate(cam, make_params(sqrt_spp=q), None, solo.data_ptr(), 0)
        out_solo = torch.empty_like(acc)
        r.finalize(solo.data_ptr(), None, out_solo.data_ptr(), w, h, q * q, True, 0)
        torch.cuda.synchronize()
        d1 = float((peer.out_t - out_nccl).abs().max())
        d2 = float((peer.out_t - out_solo).abs().max())
        line.update(max_abs_diff_peer_vs_nccl=d1, max_abs_diff_peer_vs_single_gpu=d2, mean=float(peer.out_t[..., :3].mean()))
        ok = d1 < 1e-5 and d2 < 1e-4 and float(peer.out_t[..., 3].min()) == 1.0
        line["ok"] = ok
        print(json.dumps(line), flush=True)
    dist.barrier()
    peer.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
