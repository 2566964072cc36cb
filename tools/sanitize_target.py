"""Tiny renders of every pipeline for compute-sanitizer (memcheck / racecheck / initcheck)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import raytracing_rust_b200 as rtb
pipes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 2, 3]
for name, nx, ny, ns in (("cornell_smoke_book", 37, 23, 6), ("final_scene_lit", 40, 24, 4), ("random_scene", 33, 17, 4)):
    s, w, cam = rtb.Scene.named(name, nx, ny)
    for p in pipes:
        a, r8, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(pipeline=p, background=1), want_stats=True)
        print(name, "pipe", p, "mean", float(a.mean()) / ns, "segments", st["segments"], flush=True)
rays = rtb.primary_rays(cam, nx, ny)
print("closest hit", int((s.closest_hit(w, rays)[0] >= 0).sum()))
