import sys
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np
from test_gpu_parity2 import feature_scene
from oracle_binding import OracleWorld
from raysnail_b200 import make_params, Renderer, RS_FLAG_PIXEL_CENTRE
s=feature_scene("generic")
r=Renderer(0); r.upload(s); o=OracleWorld(s)
cam=s.camera(192,120)
for flags in (RS_FLAG_PIXEL_CENTRE,0):
  for sample in (0,7):
    p=make_params(samples=9, flags=flags)
    g=r.debug_primary(cam,p,sample); oo=o.debug_primary(cam,p,sample)
    hit=oo['obj_id']>=0
    bad=np.flatnonzero(hit&(g['front']!=oo['front']))
    print(flags,sample,'bad',bad.size,'ids',np.unique(oo['obj_id'][bad]))
    for b in bad[:6]:
        print(' px',b,'id',oo['obj_id'][b],'t',g['t'][b],oo['t'][b],'front',g['front'][b],oo['front'][b],'n',g['normal'][b],oo['normal'][b],'mat',g['material'][b],oo['material'][b])
d=s.desc.contents
print('roots',[ (i,d.roots[i], d.nodes[d.roots[i]].type) for i in range(d.n_roots)])
