"""debug: list mode vs tree on one scene — image diffs and oracle check"""
import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np
from conftest import load_scene
from oracle_binding import OracleWorld
from raysnail_b200 import make_params, Renderer
name = sys.argv[1] if len(sys.argv) > 1 else "csg"
s = load_scene(name)
r = Renderer(0)
cam = s.camera(192, 120)
p = make_params(samples=9)
imgs = {}
for label, lm, w in (("list", 12, 2), ("tree2", 0, 2), ("tree4", 0, 4)):
    r.set_list_max(lm); r.set_bvh_width(w); r.upload(s)
    imgs[label] = r.render(cam, p)
o, _ = OracleWorld(s).render(cam, p)
for k, v in imgs.items():
    d = np.flatnonzero((v != o).any(axis=-1).ravel())
    print(name, k, 'pixels differing from oracle', d.size, 'max abs', float(np.abs(v - o).max()))
    for b in d[:5]:
        print('   px', b, divmod(int(b), 192), v.reshape(-1, 4)[b], o.reshape(-1, 4)[b])
for k in ("tree2", "tree4"):
    d = np.flatnonzero((imgs[k] != imgs["list"]).any(axis=-1).ravel())
    print(name, 'list vs', k, 'differ', d.size)
    for b in d[:8]:
        print('   px', b, divmod(int(b), 192), imgs["list"].reshape(-1, 4)[b], imgs[k].reshape(-1, 4)[b], o.reshape(-1, 4)[b])
