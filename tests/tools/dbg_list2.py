"""debug: find the first bounce where list mode and the tree disagree, then probe that ray"""
import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np
from conftest import load_scene
from oracle_binding import OracleWorld
from raysnail_b200 import make_params, Renderer, CameraBuilder, RS_FLAG_PIXEL_CENTRE
name = sys.argv[1] if len(sys.argv) > 1 else "csg"
s = load_scene(name)
r = Renderer(0)
ow = OracleWorld(s)
W, H = 192, 120
cam = s.camera(W, H)
def both(fn):
    out = {}
    for label, lm in (("list", 12), ("tree", 0)):
        r.set_list_max(lm); r.upload(s)
        out[label] = fn()
    return out
for depth in (1, 2, 3, 4, 5, 6, 8):
    p = make_params(sqrt_spp=3, depth=depth)
    im = both(lambda: r.render(cam, p))
    o, _ = ow.render(cam, p)
    d = np.flatnonzero((im["list"] != im["tree"]).any(axis=-1).ravel())
    print('depth', depth, 'list vs tree differ', d.size, 'list vs oracle >1e-6', int((np.abs(im["list"] - o).max(axis=-1) > 1e-6).sum()),
          'tree vs oracle >1e-6', int((np.abs(im["tree"] - o).max(axis=-1) > 1e-6).sum()))
    if d.size:
        break
p = make_params(sqrt_spp=3, depth=depth)
for b in d[:4]:
  for si in range(9):
    sc = both(lambda: r.debug_scatter(cam, p, si))
    osc = ow.debug_scatter(cam, p, si)
    pr = both(lambda: r.debug_primary(cam, p, si))
    print('sample', si)
    print('px', b, 'primary id', pr["list"]["obj_id"][b], pr["tree"]["obj_id"][b], 'kind', sc["list"]["kind"][b], sc["tree"]["kind"][b], osc["kind"][b])
    nxt = sc["list"]["next"][b]
    print('  next ray', nxt, 'tree next', sc["tree"]["next"][b])
    o3, d3 = nxt[:3], nxt[3:]
    pc = CameraBuilder().look_from(o3).look_at(o3 + d3).vup((0.123, 1, 0.0456)).fov(1.0).aperture(0.0).focus(1.0).width(1).height(1).build()
    pp = make_params(sqrt_spp=1, depth=1, flags=RS_FLAG_PIXEL_CENTRE)
    g = both(lambda: r.debug_primary(pc, pp, 0))
    oo = ow.debug_primary(pc, pp, 0)
    for k in ("obj_id", "t", "front", "normal", "point"):
        print('   ', k, g["list"][k][0], g["tree"][k][0], oo[k][0])
