import sys
sys.path.insert(0,'/root/repo')
import raytracing_rust_b200 as rtb
for nx,ny in ((1920,1080),(3840,2160)):
    s,w,cam=rtb.Scene.named("final_scene",nx,ny)
    o=rtb.default_opts()
    s.render(w,cam,nx,ny,2,o,want_sum=False,want_rgb8=False)
    for ns in (4,8,16,32,64,128):
        best=1e9
        for _ in range(2):
            _,_,st=s.render(w,cam,nx,ny,ns,o,want_sum=False,want_rgb8=False,want_stats=True)
            best=min(best,st["device_ms"])
        print(f"{nx}x{ny} spp {ns:4d}: {best:8.2f} ms {st['paths']/best/1e3:8.1f} Mpaths/s launches {st['kernel_launches']}",flush=True)
