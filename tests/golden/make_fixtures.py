#!/usr/bin/env python
"""Regenerates the committed fixtures under tests/golden/ (run in the build container, where
/root/reference exists).  1) copies the reference's scene-description data files; 2) writes
small golden vectors produced by the CPU oracle so that later oracle edits are detected."""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def copy_sdl():
    src = "/root/reference/sdl"
    dst = os.path.join(HERE, "sdl")
    os.makedirs(dst, exist_ok=True)
    for f in sorted(os.listdir(src)):
        if f.endswith(".sdl"):
            shutil.copyfile(os.path.join(src, f), os.path.join(dst, f))


def copy_assets():
    """The reference's two binary assets: the earth texture of `all_feature_scene`
    (examples/common/scene.rs:425-431) and the one image the reference's authors rendered with it
    (examples/sdl_quadrics.jpg = sdl/quadric.sdl at 800x500), used as an image-level pin of the oracle."""
    for f in ("earth-map.png", "sdl_quadrics.jpg"):
        shutil.copyfile(os.path.join("/root/reference/examples", f), os.path.join(HERE, f))


def oracle_vectors():
    import numpy as np
    from conftest import load_scene
    from oracle_binding import OracleWorld
    from raysnail_b200 import make_params, RS_FLAG_PIXEL_CENTRE
    out = {}
    for name in ["example", "quadric", "csg", "declares", "transforms", "rtow_final", "cornell", "mesh"]:
        s = load_scene(name)
        w = OracleWorld(s)
        cam = s.camera(48, 30)
        p = make_params(samples=4, flags=RS_FLAG_PIXEL_CENTRE)
        o = w.debug_primary(cam, p, 0)
        out[name + "_id"] = o["obj_id"].astype(np.int32)
        out[name + "_t"] = o["t"]
        out[name + "_n"] = o["normal"]
        p2 = make_params(samples=4)
        sc = w.debug_scatter(cam, p2, 1)
        out[name + "_kind"] = sc["kind"].astype(np.int8)
        out[name + "_next"] = sc["next"]
    np.savez_compressed(os.path.join(HERE, "oracle_vectors.npz"), **out)


def raymarching_vectors():
    import numpy as np
    from conftest import load_scene
    from oracle_binding import OracleWorld
    from raysnail_b200 import make_params, RS_FLAG_PIXEL_CENTRE
    s = load_scene("raymarching")
    o = OracleWorld(s).debug_primary(s.camera(48, 30), make_params(samples=4, flags=RS_FLAG_PIXEL_CENTRE), 0)
    np.savez_compressed(os.path.join(HERE, "raymarching_vectors.npz"), id=o["obj_id"].astype(np.int32), t=o["t"], n=o["normal"])


if __name__ == "__main__":
    copy_sdl()
    copy_assets()
    if "--only-copies" in sys.argv:
        sys.exit(0)
    if "--only-raymarching" not in sys.argv:
        oracle_vectors()
    raymarching_vectors()
    print("fixtures written")
