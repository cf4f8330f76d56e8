import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, caveengine_b200 as cb
from caveengine_b200 import scenes
from oracle import oracle as O
orc=O.port()
for name,sc,mb in (("tiny",scenes.tiny_scene(),5),("cornell",scenes.cornell_scene(),5)):
    w,h,fr=96,54,256
    want,st=orc.render(orc.assemble(sc),sc.camera,w,h,0,fr,mb)
    with cb.PathTracer(w,h,mb) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE); pt.upload_scene(sc); pt.set_camera(sc.camera)
        pt.render(0,fr,sync=True); got=pt.read_accum()
    g,r=got[...,:3].astype(np.float64),want[...,:3].astype(np.float64); m=r.mean()
    print(name,"rel rmse",np.sqrt(((g-r)**2).mean())/m,"bias",abs(g.mean()-m)/m,"far",(np.abs(g-r)>0.05*(np.abs(r)+1e-3)).any(-1).mean(),"exact px",(g==r).all(-1).mean())
