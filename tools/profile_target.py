"""Tiny, fixed workload for ncu: one scene, one pipeline, a few launches."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb

name = sys.argv[1] if len(sys.argv) > 1 else "final_scene"
nx, ny, ns, pipe = (int(x) for x in (sys.argv[2:6] if len(sys.argv) > 5 else (1920, 1080, 8, 1)))
s, w, cam = rtb.Scene.named(name, nx, ny)
o = rtb.default_opts(pipeline=pipe)
if len(sys.argv) > 6 and sys.argv[6] in ("bvh2", "bvh4", "flat"):
    o.reserved[6] = 1 if sys.argv[6] == "flat" else int(sys.argv[6][3])  # force the flat scan / the two-child / the four-wide tree
if len(sys.argv) > 7:
    o.reserved[3] = int(sys.argv[7])  # blocks of 256 threads per SM (occupancy point)
for _ in range(2):
    _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
print(name, nx, ny, ns, pipe, f"{st['device_ms']:.2f} ms", f"{st['paths']/st['device_ms']/1e3:.1f} Mpaths/s", st["kernel_launches"])
