"""Host full-sweep SAH vs device LBVH (csrc/bvh_build.cu): time of rtb_scene_upload (flatten + build + H2D) and the
render throughput on the resulting tree."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import raytracing_rust_b200 as rtb


def cloud(n, seed=1):
    rng = np.random.default_rng(seed)
    s = rtb.Scene()
    mats = [s.lambertian(s.solid(*rng.uniform(0.2, 0.9, 3))) for _ in range(8)] + [s.metal(s.solid(0.8, 0.8, 0.9), 0.1), s.dielectric(1.5)]
    side = 100.0 * (n / 1e5) ** (1 / 3)
    c = rng.uniform(-side, side, (n, 3))
    r = rng.uniform(0.3, 1.2, n)
    objs = [s.sphere(c[i], r[i], mats[i % len(mats)]) for i in range(n)]
    w = s.list(objs)
    cam = rtb.camera((0, 0, -3.2 * side), (0, 0, 0), (0, 1, 0), 40.0, 16 / 9, 0.0, 10.0, 0.0, 1.0)
    return s, w, cam


cases = [("final_scene", None)] + [(f"sphere cloud {n}", n) for n in (int(x) for x in (sys.argv[1:] or ["100000", "1000000"]))]
for label, n in cases:
    nx, ny, ns = 1920, 1080, 8
    if n is None:
        s, w, cam = rtb.Scene.named("final_scene", nx, ny)
    else:
        s, w, cam = cloud(n)
    for name, b in (("host SAH", rtb.BVH_HOST_SAH), ("device LBVH", rtb.BVH_DEVICE_LBVH)):
        if b == rtb.BVH_HOST_SAH and n is not None and n > 300000:
            print(f"{label:22s} {name:12s} skipped (host sweep is O(n log^2 n))", flush=True)
            continue
        s.set_bvh_builder(b)
        t0 = time.perf_counter()
        s.upload(w, 0)
        t_up = time.perf_counter() - t0
        s.render(w, cam, nx, ny, 1, rtb.default_opts(background=1), want_sum=False, want_rgb8=False)
        _, _, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(background=1, collect_stats=1), want_sum=False, want_rgb8=False, want_stats=True)
        _, _, st2 = s.render(w, cam, nx, ny, ns, rtb.default_opts(background=1), want_sum=False, want_rgb8=False, want_stats=True)
        print(f"{label:22s} {name:12s} upload {1e3*t_up:9.1f} ms   render {st2['paths']/st2['device_ms']/1e3:8.1f} Mpaths/s   "
              f"node tests/segment {st['node_tests']/st['segments']:6.2f}  prim tests/segment {st['prim_tests']/st['segments']:5.2f}", flush=True)
