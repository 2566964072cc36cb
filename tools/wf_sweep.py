import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
name, nx, ny, ns = "final_scene", 1920, 1080, 16
s, w, cam = rtb.Scene.named(name, nx, ny)
for cap in [int(x) for x in sys.argv[1:]]:
    o = rtb.default_opts(pipeline=2)
    o.reserved[4] = cap
    s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
    _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
    print(f"capacity {cap:9d}: {st['device_ms']:8.2f} ms {st['paths']/st['device_ms']/1e3:8.1f} Mpaths/s launches {st['kernel_launches']}", flush=True)
