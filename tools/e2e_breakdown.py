"""Where the host-buffer call (rtb_render) spends its time: deserialize, flatten+upload, render+D2H."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
nx, ny, ns = 3840, 2160, int(sys.argv[1]) if len(sys.argv) > 1 else 32
s, w, cam = rtb.Scene.named("final_scene", nx, ny)
s.render(w, cam, 64, 64, 1)
blob = s.serialize()
for want_sum in (True, False):
    for rep in range(3):
        t0 = time.perf_counter(); s2 = rtb.Scene.deserialize(blob); t1 = time.perf_counter()
        s2.upload(w, 0); t2 = time.perf_counter()
        g, g8, st = s2.render(w, cam, nx, ny, ns, rtb.default_opts(), want_sum=want_sum, want_stats=True); t3 = time.perf_counter()
        print(f"want_sum={want_sum} deserialize {1e3*(t1-t0):.1f} ms  flatten+upload {1e3*(t2-t1):.1f} ms  render call {1e3*(t3-t2):.1f} ms "
              f"(device {st['device_ms']:.1f} ms)  total {1e3*(t3-t0):.1f} ms -> {nx*ny*ns/(t3-t0)/1e6:.0f} Mpaths/s", flush=True)
