#!/usr/bin/env python
"""bench.py — headline benchmark of the raysnail hot path on B200.

Workload (BASELINE.json configs[1]): the book-1 final scene of examples/rtow_13_1.rs
(seed 7, ~480 spheres, BVH) at 1920x1080, 1024 spp (sqrt_spp 32), depth 8.
One "step" = one full render of that image (2.12 G camera paths).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1 is launched by torchrun (one rank per GPU, NCCL): the stratum rows are split across the
ranks (fixed total work = strong scaling), the per-GPU float4 accumulation buffers are
summed onto rank 0 with one NCCL reduce, rank 0 finalizes.  Rank 0 prints ONE JSON line.

`--impl reference` times the reference's CPU implementation of the same path.  The Rust
reference cannot be built in this environment (no cargo/rustc), so that arm runs the C++
restatement under oracle/ ("kind": "port") with all host threads on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WIDTH, HEIGHT, SQRT_SPP, DEPTH, SCENE_SEED = 1920, 1080, 32, 8, 7
WORKLOAD = "rtow_13_1 book-1 final scene (seed 7, 477 spheres + light, BVH), 1920x1080, 1024 spp, depth 8"
METRIC = "path samples/sec"
UNIT = "samples/s"

# algorithmic bytes per ray segment of the wavefront: 160 B/segment split by
# kernel (extend: ray read 32 + hit write 16; shade: hit read 16 + state r/w 64 + ray write 32)
BYTES_PER_SEGMENT = {"extend": 48.0, "shade": 112.0}
BYTES_PER_SAMPLE = 32.0


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of one launch of the dominant kernel, from the
    `ncu --set full` capture profiles/capture.sh made (profiles/traffic.json, written by profiles/ncu_metrics.py with
    the commit it ran and the paths of the captured wave): the bounce-0 launch of one wave, i.e. the largest launch —
    to be compared with THAT launch's algorithmic bytes (roofline.traffic_launch), not with the average over all
    bounces in algorithmic_bytes_per_launch.  Returns (bytes, description) or (None, None)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            e = json.load(open(p))[kernel + "_kernel"]
            paths = float(e.get("paths_per_wave") or 8294400)
            return float(e["per_launch_bytes"][0]), {
                "what": f"bounce 0 of one wave of {int(paths)} paths (profiles/traffic.json: {e.get('source')}, commit {e.get('commit', 'r01')})",
                "algorithmic_bytes": BYTES_PER_SEGMENT[kernel] * paths}
        except Exception:
            return None, None
    return None, None


# ---------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle (C++ restatement of the reference's CPU path)
# ---------------------------------------------------------------------------------------------
def cpu_render_sample(sqrt_spp_sample, steps=1, warmup=0):
    """Times the CPU path on the same scene/resolution with a reduced sqrt_spp (samples/s is
    spp-independent).  Thread partition = painter.rs: hardware threads + 1, interleaved rows."""
    from oracle_binding import OracleWorld
    from raysnail_b200 import Scene, make_params
    scene = Scene.builtin("rtow_final", SCENE_SEED)
    world = OracleWorld(scene)
    cam = scene.camera(WIDTH, HEIGHT)
    params = make_params(sqrt_spp=sqrt_spp_sample, depth=DEPTH)
    cores = os.cpu_count() or 1
    times, samples, segments = [], 0, 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        _, st = world.render(cam, params, threads=cores + 1)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            samples, segments = int(st.n_samples), int(st.n_segments)
    total = sum(times)
    return {"samples_per_s": samples * len(times) / total, "mrays_per_s": segments * len(times) / total / 1e6, "ms_per_step": 1e3 * total / len(times),
            "cores": cores, "sample": f"same scene and 1920x1080 frame at {sqrt_spp_sample * sqrt_spp_sample} spp ({samples} paths per step)"}


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    r = cpu_render_sample(1, steps=max(args.steps, 1), warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["samples_per_s"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "mrays_per_s": r["mrays_per_s"],
        "config": {"workload": WORKLOAD, "note": "CPU arm: C++ restatement of the Rust reference (no Rust toolchain here); bounded sample per step"},
        "cpu_baseline": {"value": r["samples_per_s"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
        "e2e": {"value": r["samples_per_s"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def scene_bytes(desc):
    import ctypes as C
    from raysnail_b200 import _abi
    d = desc.contents
    return (d.n_nodes * C.sizeof(_abi.RsNode) + d.n_roots * (4 + 48) + d.n_lights * 4 + d.n_xforms * C.sizeof(_abi.RsXform) + d.n_tris * (72 + 72 + 4) +
            d.n_materials * C.sizeof(_abi.RsMaterial) + d.n_textures * C.sizeof(_abi.RsTexture) + d.n_texel_bytes + d.n_perlin_perm * 4 + d.n_perlin_values * 8)


E2E_WARM = ("warm: the world is unchanged between steps, so rs_upload_scene re-uploads the scene tables but keeps the SAH tree "
            "and its device copy (tree cache); wave buffers already allocated")
E2E_COLD = ("cold: a fresh context per step — rs_create + rs_upload_scene (SAH build from nothing, tree upload) + rs_render with "
            "host mask and host image, wave buffers allocated inside the timed region; the first frame a caller sees")


def e2e_cold(device, scene, cam, params, host_mask_ptr, host_img_ptr, samples_per_step, bvh_width=2, wave_streams=4, reps=2):
    """First-frame e2e: nothing cached.  Every repetition creates its own context, so neither the tree cache nor the
    wave buffers of an earlier call can help; wall clock around create + upload + render + destroy."""
    from raysnail_b200 import Renderer
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        r = Renderer(device)
        try:
            r.set_bvh_width(bvh_width)
            r.set_wave_streams(wave_streams)
            r.upload(scene)
            r.render_to_ptr(cam, params, host_mask_ptr, host_img_ptr)
        finally:
            r.close()
        times.append(time.perf_counter() - t0)
    dt = min(times)  # (the first repetition may also pay one-time driver work that is no property of the frame)
    return {"value": samples_per_step / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "ms_per_step_all": [t * 1e3 for t in times], "kind": E2E_COLD}


DEFAULT_WAVE_PATHS = 1 << 26  # csrc/backend.cu render_into: paths in flight per wave when rs_render_params.max_wave_paths is 0


def load_bench_scene(name):
    """rtow_final (the headline workload) or another BASELINE.json config: an .sdl file name under
    tests/golden/sdl (example, quadric, csg, declares, transforms), cornell, mesh:<grid>, mesh_blinn:<grid>."""
    from raysnail_b200 import Scene
    sdl = os.path.join(ROOT, "tests", "golden", "sdl", name + ".sdl")
    if os.path.exists(sdl):
        return Scene.from_sdl(sdl)
    if name.startswith("mesh"):
        kind, _, grid = name.partition(":")
        return Scene.builtin(kind, SCENE_SEED, int(grid or 708))
    return Scene.builtin(name, SCENE_SEED)


# BASELINE.json's other configs (driver-visible numbers next to the headline line): name, scene, frame, sqrt(spp)
OTHER_CONFIGS_1GPU = [
    ("C1 sdl/example.sdl 800x500 64 spp (--samples 65 -> 8x8 strata)", "example", 800, 500, 8),
    ("C3a sdl/quadric.sdl 1920x1080 256 spp", "quadric", 1920, 1080, 16),
    ("C3b sdl/csg.sdl 1920x1080 256 spp", "csg", 1920, 1080, 16),
    ("C4 1M-triangle mesh (mesh_blinn:708, 1 001 112 triangles), Blinn-Phong, 2048x2048 64 spp", "mesh_blinn:708", 2048, 2048, 8),
    ("C5 Cornell box 3840x2160 256 spp (the 4096 spp of the config on N > 1)", "cornell", 3840, 2160, 16),
]
OTHER_CONFIGS_NGPU = [("C5 Cornell box 3840x2160 4096 spp", "cornell", 3840, 2160, 64)]


def measure_config(renderer, dev, label, scene_name, width, height, q, steps, world_size, rank, peak):
    """One BASELINE.json config through the same sharded step as the headline: `steps` timed frames (CUDA events,
    barrier + synchronize on both sides, max over ranks) after one warm-up frame, then one frame with per-kernel
    event timing for the breakdown.  All ranks call this; every rank returns the same dict."""
    import torch
    import torch.distributed as dist
    from raysnail_b200 import RS_FLAG_KERNEL_TIMING, make_params
    from raysnail_b200.distributed import PeerImage, stratum_rows
    scene = load_bench_scene(scene_name)
    renderer.upload(scene)
    cam = scene.camera(width, height)
    stream = torch.cuda.current_stream().cuda_stream
    r0, r1 = stratum_rows(rank, world_size, q)
    params = make_params(sqrt_spp=q, depth=DEPTH, rows=(r0, r1))
    params_t = make_params(sqrt_spp=q, depth=DEPTH, rows=(r0, r1), flags=RS_FLAG_KERNEL_TIMING)
    peer = PeerImage(renderer, width, height) if world_size > 1 else None
    if peer is not None:
        accum, out = peer.accum_t, peer.out_t
    else:
        accum = torch.zeros((height, width, 4), dtype=torch.float32, device=dev)
        out = torch.empty_like(accum)

    def step(p):
        accum.zero_()
        renderer.render_accumulate(cam, p, None, accum.data_ptr(), stream)
        if peer is not None:
            peer.reduce_finalize(q * q, True, stream, None)
        else:
            renderer.finalize(accum.data_ptr(), None, out.data_ptr(), width, height, q * q, True, stream)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step(params)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        step(params)
    ev1.record()
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    st = renderer.stats()
    seg = torch.tensor([float(st.n_segments), float(st.n_samples)], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(seg, op=dist.ReduceOp.SUM)
    ms_step = float(ms.item()) / steps
    n_segments, n_samples = [float(x) for x in seg.tolist()]
    step(params_t)
    barrier()
    kt = renderer.stats()
    shares = {"extend": kt.ms_extend, "shade": kt.ms_shade, "raygen": kt.ms_raygen, "accumulate": kt.ms_accumulate}
    dom = max(("extend", "shade"), key=lambda k: shares[k])
    # the dominant kernel's algorithmic bytes over its own summed launch time (rank 0's share of the segments)
    achieved = BYTES_PER_SEGMENT[dom] * float(kt.n_segments) / (shares[dom] * 1e-3) / 1e9 if shares[dom] > 0 else 0.0
    e2e = None
    if world_size == 1:
        # the reference-facing call with HOST buffers: rs_upload_scene + rs_render (host mask in, host image out)
        host_img = torch.empty((height, width, 4), dtype=torch.float32).pin_memory()
        host_mask = torch.ones((height, width), dtype=torch.uint8).pin_memory()
        p_all = make_params(sqrt_spp=q, depth=DEPTH)

        def e2e_step():
            renderer.upload(scene)
            renderer.render_to_ptr(cam, p_all, host_mask.data_ptr(), host_img.data_ptr())

        e2e_step()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        dt = (time.perf_counter() - t0) / steps
        e2e = {"value": width * height * q * q / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": int(scene_bytes(scene.desc) + height * width),
               "d2h_bytes_per_step": height * width * 16, "kind": E2E_WARM,
               "cold": e2e_cold(dev.index, scene, cam, p_all, host_mask.data_ptr(), host_img.data_ptr(), width * height * q * q, reps=1)}
    if peer is not None:
        barrier()
        peer.close()
    return {"workload": label, "e2e": e2e, "value": n_samples / (ms_step / 1e3), "unit": UNIT, "mrays_per_s": n_segments / (ms_step / 1e3) / 1e6, "ms_per_step": ms_step,
            "steps": steps, "segments_per_sample": n_segments / max(n_samples, 1.0),
            "roofline": {"bound": "hbm", "kernel": dom + "_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "kernel_ms_per_step": shares}}


def run_gpu(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from raysnail_b200 import RS_FLAG_KERNEL_TIMING, Renderer, Scene, make_params
    from raysnail_b200.distributed import PeerImage, reduce_accum, stratum_rows

    world_size, rank, local_rank = env_int("WORLD_SIZE", 1), env_int("RANK", 0), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world_size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    width, height, q, depth = args.width, args.height, args.sqrt_spp, DEPTH
    scene = load_bench_scene(args.scene)
    renderer = Renderer(local_rank)
    renderer.set_bvh_width(args.bvh_width)
    renderer.set_wave_streams(args.wave_streams)
    renderer.upload(scene)
    cam = scene.camera(width, height)
    stream = torch.cuda.current_stream().cuda_stream

    # strata rows of this rank: contiguous, balanced
    r0, r1 = stratum_rows(rank, world_size, q)
    params = make_params(sqrt_spp=q, depth=depth, rows=(r0, r1), max_wave_paths=args.wave)
    params_t = make_params(sqrt_spp=q, depth=depth, rows=(r0, r1), max_wave_paths=args.wave, flags=RS_FLAG_KERNEL_TIMING)
    # N > 1: the per-rank accumulators are NVLink peer memory and the combine step is the fused
    # reduce + finalize kernel (rs_finalize_peers); `--reduce nccl` times the library baseline
    # (ncclReduce + rs_finalize on rank 0) instead
    peer = PeerImage(renderer, width, height) if (world_size > 1 and args.reduce == "peer") else None
    if peer is not None:
        accum, out = peer.accum_t, peer.out_t
    else:
        accum = torch.zeros((height, width, 4), dtype=torch.float32, device=dev)
        out = torch.empty_like(accum)

    def combine(d_mask=None):
        if peer is not None:
            peer.reduce_finalize(q * q, True, stream, d_mask)
            return
        reduce_accum(accum, 0)
        if rank == 0:
            renderer.finalize(accum.data_ptr(), d_mask, out.data_ptr(), width, height, q * q, True, stream)

    def step(p):
        accum.zero_()
        renderer.render_accumulate(cam, p, None, accum.data_ptr(), stream)
        combine()

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(params)
    barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step(params)
    ev1.record()
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    st = renderer.stats()
    seg = torch.tensor([float(st.n_segments), float(st.n_samples), float(st.n_launches)], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(seg, op=dist.ReduceOp.SUM)
    n_segments, n_samples, n_launches = [float(x) for x in seg.tolist()]

    # per-kernel CUDA-event timing (events on the launching stream), one extra step
    barrier()
    step(params_t)
    barrier()
    kt = renderer.stats()
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e through the public API with host buffers -------------------------------------
    host_img = torch.empty((height, width, 4), dtype=torch.float32).pin_memory()
    host_mask = torch.ones((height, width), dtype=torch.uint8).pin_memory()
    d_mask = torch.empty((height, width), dtype=torch.uint8, device=dev)
    h2d = scene_bytes(scene.desc) + height * width + C.sizeof(type(cam)) + C.sizeof(type(params))
    d2h = height * width * 16 if rank == 0 else 0

    def e2e_step():
        if world_size == 1:
            # the reference-facing call itself: rs_upload_scene (host scene buffers -> HBM) + rs_render with a HOST
            # pixel mask and a HOST output image (the C ABI's shot_to_target).  WARM: the world does not change between
            # steps, so the upload keeps the SAH tree (tree cache, filled by the untimed first call) and the all-ones
            # mask is recognised as "no mask"; the cold first frame is measured separately (e2e_cold)
            renderer.upload(scene)
            renderer.render_to_ptr(cam, params, host_mask.data_ptr(), host_img.data_ptr())
            return
        renderer.upload(scene)                       # host scene buffers -> HBM (incl. BVH build)
        d_mask.copy_(host_mask, non_blocking=True)   # pixel mask from pinned host memory
        # a mask that masks nothing is passed as NULL, as rs_render / rs_multi_render do with their host masks
        # (memchr; backend.cu): the unmasked path generates the camera rays inside the first extend
        mask_ptr = d_mask.data_ptr() if not bool(host_mask.all()) else None
        accum.zero_()
        renderer.render_accumulate(cam, params, mask_ptr, accum.data_ptr(), stream)
        combine(mask_ptr)
        if rank == 0:
            host_img.copy_(out, non_blocking=True)   # the image the caller gets back
        torch.cuda.synchronize()

    e2e_steps = max(1, min(args.steps, 3))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    total_samples_per_step = float(width * height * q * q)
    e2e_value = total_samples_per_step * e2e_steps / (float(e2e_ms.item()) / 1e3)
    cold = None
    if world_size == 1:
        p_cold = make_params(sqrt_spp=q, depth=depth, max_wave_paths=args.wave)
        cold = e2e_cold(local_rank, scene, cam, p_cold, host_mask.data_ptr(), host_img.data_ptr(), total_samples_per_step, args.bvh_width, args.wave_streams)

    # the other BASELINE.json configs, outside the timed headline region (all ranks take part; N = 1: C1, C3a, C3b, C4, C5
    # at 256 spp; N > 1: C5 at its full 4096 spp)
    others, others_clocks = [], None
    if not (args.no_cpu or args.no_others) and (args.scene, width, height, q) == ("rtow_final", WIDTH, HEIGHT, SQRT_SPP):
        peak0 = measured_peak_gbs()[0]
        osampler = ClockSampler(local_rank)
        if rank == 0:
            osampler.start()
        for label, name, w_, h_, q_ in (OTHER_CONFIGS_1GPU if world_size == 1 else OTHER_CONFIGS_NGPU):
            try:
                others.append(measure_config(renderer, dev, label, name, w_, h_, q_, 2, world_size, rank, peak0))
            except Exception as exc:  # a config that cannot run is reported, not hidden
                others.append({"workload": label, "error": f"{type(exc).__name__}: {exc}"})
        if rank == 0:
            others_clocks = osampler.stop()

    # ---- N > 1: is the N-GPU image the one-GPU image, and what does the ONE-call C ABI (rs_multi_render) deliver? ----
    image_check, multi_call = None, None
    if world_size > 1:
        cpu_group = dist.new_group(backend="gloo")  # the idle ranks wait on the host: an NCCL barrier would spin on their GPUs
        renderer.upload(scene)  # (the other configs above left their own scene in the context)
        step(params)
        barrier()
        if rank == 0:
            sharded = out.clone()
            # the whole frame on rank 0's GPU alone, same seed: every pixel is the same set of per-sample values, summed in
            # N groups instead of one, so the images may differ by the f32 rounding of that grouping only
            accum1 = torch.zeros((height, width, 4), dtype=torch.float32, device=dev)
            out1 = torch.empty_like(accum1)
            renderer.render_accumulate(cam, make_params(sqrt_spp=q, depth=depth, max_wave_paths=args.wave), None, accum1.data_ptr(), stream)
            renderer.finalize(accum1.data_ptr(), None, out1.data_ptr(), width, height, q * q, True, stream)
            torch.cuda.synchronize()
            diff = (sharded - out1).abs()
            image_check = {"vs": "the whole frame rendered on rank 0's GPU alone (same Philox key)", "max_abs_diff": float(diff.max().item()),
                           "mean_abs_diff": float(diff.mean().item()), "mean_value": float(out1[..., :3].mean().item()),
                           "ok": bool(torch.allclose(sharded, out1, rtol=2e-5, atol=2e-6))}
            del accum1, out1, diff
            # rs_create_multi / rs_multi_render: what a single-process host (the Rust crate) binds — host scene in, host image
            # out, all N devices behind one call, no torch.distributed anywhere on the path
            try:
                from raysnail_b200 import MultiRenderer
                mr = MultiRenderer(list(range(world_size)))
                try:
                    mr.upload(scene)
                    p_all = make_params(sqrt_spp=q, depth=depth, max_wave_paths=args.wave)
                    mr.render_to_ptr(cam, p_all, None, host_img.data_ptr())
                    m_steps = max(1, min(args.steps, 3))
                    t0 = time.perf_counter()
                    for _ in range(m_steps):
                        mr.render_to_ptr(cam, p_all, None, host_img.data_ptr())
                    dt = (time.perf_counter() - t0) / m_steps
                    got, want = host_img.clone(), sharded.cpu()
                    same = bool(torch.allclose(got, want, rtol=2e-5, atol=2e-6))
                    d = (got - want).abs()
                    bad = d > (2e-6 + 2e-5 * want.abs())
                    diag = {"max_abs_diff": float(d.max().item()), "pixels_off": int(bad.any(dim=-1).sum().item()), "nan": int(torch.isnan(got).sum().item()),
                            "first_off": [int(v) for v in bad.any(dim=-1).nonzero()[0].tolist()] if bool(bad.any()) else None}
                    multi_call = {"call": "rs_multi_render (one process, one host thread per device, host image out)", "value": total_samples_per_step / dt,
                                  "unit": UNIT, "ms_per_step": dt * 1e3, "steps": m_steps, "devices": world_size, "image_equals_torchrun_image": same, "image_diff": diag,
                                  "timing": "host wall clock around the call (it returns after the image is in host memory)"}
                finally:
                    mr.close()
            except Exception as exc:
                multi_call = {"error": f"{type(exc).__name__}: {exc}"}
        dist.barrier(group=cpu_group)

    if rank == 0:
        value = total_samples_per_step * args.steps / (ms_total / 1e3)
        peak, peak_src = measured_peak_gbs()
        # dominant kernel by CUDA-event time on rank 0's stream
        shares = {"extend": kt.ms_extend, "shade": kt.ms_shade, "raygen": kt.ms_raygen, "accumulate": kt.ms_accumulate}
        dom = max(("extend", "shade"), key=lambda k: shares[k])
        # waves of one render (the partition of render_into, csrc/backend.cu): P pixels x K samples each
        cap = max(args.wave or DEFAULT_WAVE_PATHS, 1024)
        s_total = (r1 - r0) * q
        P = min(width * height, cap)
        K = max(1, min(s_total, cap // P))
        waves = max(1, -(-(width * height) // P) * -(-s_total // K))
        n_dom_launches = float(waves * depth)
        seg_rank0 = float(kt.n_segments)
        alg_bytes_per_launch = BYTES_PER_SEGMENT[dom] * seg_rank0 / n_dom_launches
        avg_ms = shares[dom] / n_dom_launches
        achieved = alg_bytes_per_launch / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        pipeline_bytes = 160.0 * n_segments + BYTES_PER_SAMPLE * n_samples
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD if (args.scene, width, height, q) == ("rtow_final", WIDTH, HEIGHT, SQRT_SPP)
                       else f"{args.scene} {width}x{height} {q * q} spp depth 8 (not the headline workload)",
                       "parallelism": f"strata-rows x{world_size}",
                       "oracle": "port, canonical rules (C++ restatement of the Rust reference with the tie / range rules of DESIGN.md §2; parity unpinned)",
                       "combine": ("none (single GPU: finalize only)" if world_size == 1 else
                                   ("fused reduce+finalize kernel over NVLink peer memory" if args.reduce == "peer" else "ncclReduce + finalize")), "l2": "path/ray/hit buffers of one wave (> 12 GB at 64 Mi paths) exceed the 126 MB L2; no flush needed",
                       "wave_paths": args.wave or DEFAULT_WAVE_PATHS, "wave_streams": args.wave_streams},
            "mrays_per_s": n_segments / (ms_total / args.steps / 1e3) / 1e6,
            "segments_per_sample": n_segments / max(n_samples, 1.0),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "kind": E2E_WARM, "cold": cold},
            "gpu_launches": int(n_launches * args.steps),
            "roofline": {"bound": "hbm", "kernel": dom + "_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(dom)[0],
                         "traffic_launch": ncu_traffic(dom)[1],
                         "peak_source": peak_src, "avg_launch_ms": avg_ms,
                         "algorithmic_bytes_per_launch": alg_bytes_per_launch,
                         "pipeline": {"bytes_per_step": pipeline_bytes, "achieved": pipeline_bytes / (ms_total / args.steps / 1e3) / 1e9 / world_size,
                                      "frac": pipeline_bytes / (ms_total / args.steps / 1e3) / 1e9 / world_size / peak},
                         "kernel_ms_per_step": shares},
            "clocks": clocks,
        }
        if image_check is not None:
            line["image_check"] = image_check
        if multi_call is not None:
            line["multi_call"] = multi_call
        if others:
            line["other_configs"] = others
            line["other_configs_clocks"] = others_clocks
        if world_size == 1 and not args.no_cpu:
            c = cpu_render_sample(2)
            line["cpu_baseline"] = {"value": c["samples_per_s"], "unit": UNIT, "cores": c["cores"], "kind": "port", "sample": c["sample"],
                                    "mrays_per_s": c["mrays_per_s"]}
        print(json.dumps(line), flush=True)
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--sqrt-spp", type=int, default=SQRT_SPP)
    ap.add_argument("--wave", type=int, default=0, help="max paths in flight per wave (0 = library default)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (and the other configs)")
    ap.add_argument("--no-others", action="store_true", help="skip the other BASELINE.json configs (other_configs)")
    ap.add_argument("--reduce", default="peer", choices=["peer", "nccl"],
                    help="N > 1 combine step: peer = fused reduce+finalize kernel over NVLink peer memory (default); nccl = ncclReduce + finalize")
    ap.add_argument("--bvh-width", type=int, default=2, choices=[2, 4], help="world BVH: 2 = binary (default), 4 = QBVH")
    ap.add_argument("--wave-streams", type=int, default=4, choices=[1, 2, 3, 4], help="streams consecutive waves rotate over")
    ap.add_argument("--scene", default="rtow_final", help="other BASELINE.json configs (see load_bench_scene); default = headline")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
