"""Quick per-scene throughput probe (not the bench): Mpaths/s and Msegments/s per pipeline."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import raytracing_rust_b200 as rtb

cases = [("two_spheres", 400, 225, 100), ("random_scene", 1920, 1080, 16), ("two_perlin_sphere", 1920, 1080, 16),
         ("earth", 1920, 1080, 16), ("simple_light", 1280, 720, 16), ("cornell_box", 1024, 1024, 32),
         ("cornell_smoke", 1024, 1024, 32), ("final_scene", 1920, 1080, 16), ("final_scene", 3840, 2160, 8)]
pipes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1]
for name, nx, ny, ns in cases:
    s, w, cam = rtb.Scene.named(name, nx, ny)
    for pipe in pipes:
        o = rtb.default_opts(pipeline=pipe)
        s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)  # warm-up + upload
        _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
        ms = st["device_ms"]
        print(f"{name:18s} {nx}x{ny}x{ns} pipe={pipe}: {ms:9.2f} ms  {st['paths']/ms/1e3:9.1f} Mpaths/s "
              f"{st['segments']/ms/1e3:9.1f} Mseg/s  seg/path {st['segments']/st['paths']:.2f} launches {st['kernel_launches']}",
              flush=True)
