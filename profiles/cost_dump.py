#!/usr/bin/env python
"""Per-ray traversal cost analysis (profiling build only: make VARIANT=_cost TUNE=-DRS_TUNE_COST_DUMP=1).

For one wave of the given scene it dumps, per bounce, every ray of extend with its node visits and
leaf tests, then reports the SIMT efficiency of the warps as extend groups them (sum of lane costs /
32 x sum of per-warp maxima) and what re-ordering the rays by various keys would give.
    RAYSNAIL_B200_LIB=raysnail_b200/libraysnail_b200_cost.so python profiles/cost_dump.py [scene] [w h]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def eff(cost, order=None):
    c = cost if order is None else cost[order]
    n = len(c) // 32 * 32
    g = c[:n].reshape(-1, 32)
    return float(g.sum() / max(1, 32 * g.max(axis=1).sum()))


def main():
    from bench import load_bench_scene
    from raysnail_b200 import Renderer, make_params
    scene_name = sys.argv[1] if len(sys.argv) > 1 else "rtow_final"
    w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
    scene = load_bench_scene(scene_name)
    r = Renderer(0)
    r.set_wave_streams(1)
    r.upload(scene)
    cam = scene.camera(w, h)
    for bounce in (0, 1, 2, 4):
        path = f"/tmp/cost_{bounce}.bin"
        os.environ["RAYSNAIL_B200_COST_DUMP"] = f"{path}:{bounce}"
        r.render(cam, make_params(sqrt_spp=1, max_wave_paths=w * h))
        raw = np.fromfile(path, dtype=np.uint8)
        cap, nf, nb, b = np.frombuffer(raw[:16].tobytes(), dtype=np.uint32)
        rec = np.frombuffer(raw[16:16 + cap * 64].tobytes(), dtype=np.float64).reshape(cap, 8)
        cost = np.frombuffer(raw[16 + cap * 64:16 + cap * 72].tobytes(), dtype=np.int32).reshape(cap, 2)
        print(f"== {scene_name} bounce {b}: front {nf} back {nb}")
        for name, sl in (("front", slice(0, nf)), ("back", slice(cap - nb, cap))):
            o, d, c = rec[sl, 0:3], rec[sl, 3:6], cost[sl]
            if name == "back":
                o, d, c = o[::-1], d[::-1], c[::-1]
            n = len(c)
            if n < 64:
                continue
            nodes, leaves = c[:, 0].astype(np.float64), c[:, 1].astype(np.float64)
            work = 50.0 * nodes + 110.0 * leaves  # rough instruction weights: node visit ~50, f64 object test ~110
            print(f"  {name}: n {n} nodes/ray {nodes.mean():.1f} (p50 {np.median(nodes):.0f} p99 {np.percentile(nodes, 99):.0f} max {nodes.max():.0f}) leaves/ray {leaves.mean():.2f}")
            octant = (d[:, 0] > 0).astype(int) | ((d[:, 1] > 0).astype(int) << 1) | ((d[:, 2] > 0).astype(int) << 2)
            lo, hi = np.percentile(o, 1, axis=0), np.percentile(o, 99, axis=0)
            cell = np.clip(((o - lo) / np.maximum(hi - lo, 1e-9) * 4).astype(int), 0, 3)
            cellkey = cell[:, 0] | (cell[:, 1] << 2) | (cell[:, 2] << 4)
            dyb = np.clip(((d[:, 1] + 1.0) * 4).astype(int), 0, 7)
            keys = {"octant": octant, "octant+cell": octant | (cellkey << 3), "dy8": dyb, "dy8+octant": (dyb << 3) | octant,
                    "dy8+octant+cell": (((dyb << 3) | octant) << 6) | cellkey}
            print(f"    SIMT efficiency now: nodes {eff(nodes):.3f} leaves {eff(leaves):.3f} weighted {eff(work):.3f}")
            for kn, k in keys.items():
                order = np.argsort(k, kind="stable")
                print(f"    sorted by {kn:16s}: nodes {eff(nodes, order):.3f} leaves {eff(leaves, order):.3f} weighted {eff(work, order):.3f}")
            order = np.argsort(work, kind="stable")
            print(f"    sorted by true cost (bound): weighted {eff(work, order):.3f}")
    r.close()


if __name__ == "__main__":
    main()
