import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
from conftest import load_scene
from oracle_binding import OracleWorld
from raysnail_b200 import RS_FLAG_PIXEL_CENTRE, make_params, Renderer, Scene
r = Renderer(0)
W, H = 192, 120
def compare(scene, label, flags, samples=(0,)):
    r.upload(scene)
    o_w = OracleWorld(scene)
    cam = scene.camera(W, H)
    p = make_params(samples=9, flags=flags)
    for s in samples:
        g = r.debug_primary(cam, p, s)
        g2 = r.debug_primary(cam, p, s)
        o = o_w.debug_primary(cam, p, s)
        mism = np.flatnonzero(g["obj_id"] != o["obj_id"])
        tm = np.flatnonzero((g["obj_id"] == o["obj_id"]) & (o["obj_id"] >= 0) & (np.abs(g["t"] - o["t"]) > 1e-9 * np.abs(o["t"])))
        print(label, "sample", s, "ray equal", np.array_equal(g["ray"], o["ray"]), "deterministic", np.array_equal(g["t"], g2["t"]),
              "id mism", mism.tolist(), "gpu", g["obj_id"][mism].tolist(), g["t"][mism].tolist(), "orc", o["obj_id"][mism].tolist(), o["t"][mism].tolist(),
              "t mism", tm.tolist(), g["t"][tm].tolist(), o["t"][tm].tolist())
        for m in mism.tolist() + tm.tolist():
            print("   ray", m, [float.hex(float(x)) for x in g["ray"][m]])
full = load_scene("raymarching")
compare(full, "full-centre", RS_FLAG_PIXEL_CENTRE, (0, 7))
compare(full, "full-jitter", 0, (0, 7))
# bulb only, same camera
s = Scene()
white = s.lambertian(s.solid(0.5, 0.5, 0.5))
lt = s.sphere((300, 400, 100), 12.0, s.diffuse_light(s.solid(1, 1, 1), 1.5))
s.add(lt); s.add_light(lt)
s.add(s.raymarcher(white))
s.set_camera((13.0 * 0.7, -1.7 * 0.7, 3.0 * 0.7), (0.0, -0.4, 0.0), (0, 1, 0), fov=20.0, aperture=0.01, focus=10.0)
compare(s, "bulb+light-centre", RS_FLAG_PIXEL_CENTRE, (0, 7))
compare(s, "bulb+light-jitter", 0, (0, 7))
