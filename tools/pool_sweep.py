"""Sweep the pool pipeline's knobs (opts.reserved[0..]) on one scene: device-event ms, best of 3."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
name = sys.argv[1] if len(sys.argv) > 1 else "final_scene"
nx, ny, ns = (int(x) for x in sys.argv[2:5]) if len(sys.argv) > 4 else (1920, 1080, 16)
s, w, cam = rtb.Scene.named(name, nx, ny)
for knobs in [tuple(int(x) for x in a.split(",")) for a in sys.argv[5:]] or [(16,)]:
    o = rtb.default_opts(pipeline=rtb.PIPE_POOL)
    for i, k in enumerate(knobs):
        o.reserved[i] = k
    s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
    best = 1e9
    for _ in range(3):
        _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
        best = min(best, st["device_ms"])
    print(f"{name} knobs {knobs}: {best:8.2f} ms {st['paths']/best/1e3:8.1f} Mpaths/s", flush=True)
