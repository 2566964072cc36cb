"""Bench-shape launch (final_scene 3840x2160 x 64 spp) for several item sizes (RTB200_ITEM_SAMPLES is read per launch)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
nx, ny, ns = 3840, 2160, 64
s, w, cam = rtb.Scene.named("final_scene", nx, ny)
s.render(w, cam, nx, ny, 2, want_sum=False, want_rgb8=False)
for v in sys.argv[1:]:
    os.environ["RTB200_ITEM_SAMPLES"] = v
    best = 1e9
    for _ in range(3):
        _, _, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(), want_sum=False, want_rgb8=False, want_stats=True)
        best = min(best, st["device_ms"])
    print(f"item samples {v:>4s}: {best:8.2f} ms  {st['paths'] / best / 1e3:8.1f} Mpaths/s", flush=True)
