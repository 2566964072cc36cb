import sys
sys.path.insert(0,'/root/repo')
import numpy as np
import raytracing_rust_b200 as rtb
for name,nx,ny,ns in (("cornell_box",256,256,16),("final_scene_lit",320,180,16),("two_spheres",200,112,8)):
    s,w,cam=rtb.Scene.named(name,nx,ny)
    bg = rtb.BG_SKY if name=="two_spheres" else rtb.BG_BLACK
    a,_,sa=s.render(w,cam,nx,ny,ns,rtb.default_opts(pipeline=1,background=bg),want_stats=True)
    b,_,sb=s.render(w,cam,nx,ny,ns,rtb.default_opts(pipeline=3,background=bg),want_stats=True)
    print(name,"mean mega",a.mean()/ns,"pool",b.mean()/ns,"max abs diff",np.abs(a-b).max(),"segments",sa["segments"],sb["segments"],"ms",sa["device_ms"],sb["device_ms"],flush=True)
