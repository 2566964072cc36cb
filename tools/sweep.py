"""Sweep the megakernel scheduler thresholds / occupancy target on one scene."""
import sys, itertools
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
name = sys.argv[1] if len(sys.argv) > 1 else "final_scene"
nx, ny, ns = (int(x) for x in sys.argv[2:5]) if len(sys.argv) > 4 else (1920, 1080, 16)
s, w, cam = rtb.Scene.named(name, nx, ny)
o = rtb.default_opts(collect_stats=1)
_, _, st = s.render(w, cam, nx, ny, 4, o, want_sum=False, want_rgb8=False, want_stats=True)
print(name, {k: (round(v / st["segments"], 2) if k.endswith("tests") else v) for k, v in st.items() if k in ("node_tests", "prim_tests", "medium_tests", "segments", "paths")})
grid = [tuple(int(x) for x in a.split(",")) for a in sys.argv[5:]] or [(16, 8, 8, 3)]
for knobs in grid:  # k_node, k_leaf, k_shade, occupancy point[, -, smem carveout %]
    kn, kl, ks, mb = knobs[:4]
    o = rtb.default_opts()
    for i, k in enumerate(knobs):
        o.reserved[i] = k
    s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
    best = 1e9
    for _ in range(2):
        _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
        best = min(best, st["device_ms"])
    print(f"k_node {kn:2d} k_leaf {kl:2d} k_shade {ks:2d} minblocks {mb} extra {knobs[4:]}: {best:8.2f} ms {st['paths']/best/1e3:8.1f} Mpaths/s", flush=True)
