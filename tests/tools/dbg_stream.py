"""debug: streaming extend vs batch extend on the small mesh scene"""
import os, sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np
from conftest import load_scene
from raysnail_b200 import make_params, Renderer
name = sys.argv[1] if len(sys.argv) > 1 else "mesh"
s = load_scene(name)
r = Renderer(0)
cam = s.camera(192, 120)
for depth in (2, 3, 8):
    for q in (1, 3):
        p = make_params(sqrt_spp=q, depth=depth, gamma=False)
        im = {}
        for mode in ("1", "0"):
            os.environ["RAYSNAIL_B200_STREAM"] = mode
            r.upload(s)
            im[mode] = r.render(cam, p)
            im[mode + "b"] = r.render(cam, p)
        d = np.flatnonzero((im["1"] != im["0"]).any(axis=-1).ravel())
        print('depth', depth, 'q', q, 'stream vs batch differ', d.size, 'stream vs itself', int((im["1"] != im["1b"]).any(axis=-1).sum()),
              'batch vs itself', int((im["0"] != im["0b"]).any(axis=-1).sum()))
        for b in d[:5]:
            print('   px', divmod(int(b), 192), im["1"].reshape(-1, 4)[b], im["0"].reshape(-1, 4)[b])
