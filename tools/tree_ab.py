"""A/B of the megakernel's three closest-hit forms on one build: opts.reserved[6] = 1 (flat scan) / 2 (two-child BVH) / 4 (four-wide BVH).
usage: tree_ab.py [scene ...]   (device-event ms, best of 3)"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
CASES = {"cornell_box": (1024, 1024, 32), "two_perlin_sphere": (1920, 1080, 16), "cornell_smoke": (1024, 1024, 32),
         "two_spheres": (400, 225, 100), "simple_light": (1280, 720, 16), "earth": (1920, 1080, 16), "random_scene": (480, 270, 4),
         "final_scene": (480, 270, 2)}
import numpy as np


def synthetic(n, kind, nx, ny):
    """n primitives of one kind scattered in a 20-unit cube above a ground sphere, sky-lit (where does the flat scan stop paying?)"""
    rng = np.random.default_rng(n)
    s = rtb.Scene()
    objs = [s.sphere((0, -1000, 0), 1000.0, s.lambertian(s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.9, 0.9))))]
    for i in range(n - 1):
        c = rng.uniform((-10, 0.5, -10), (10, 8, 10))
        m = [s.lambertian(s.solid(*rng.uniform(0.2, 0.9, 3))), s.metal(s.solid(0.8, 0.8, 0.8), 0.1), s.dielectric(1.5)][i % 3]
        if kind == "sphere":
            objs.append(s.sphere(tuple(c), float(rng.uniform(0.4, 1.2)), m))
        elif kind == "cube":
            h = rng.uniform(0.4, 1.2, 3)
            objs.append(s.translate(s.rotate(1, s.cube(tuple(-h), tuple(h), m), float(rng.uniform(0, 90))), tuple(c)))
        else:
            h = rng.uniform(0.5, 1.5, 2)
            objs.append(s.rect(i % 3, float(c[0] - h[0]), float(c[1] - h[1]), float(c[0] + h[0]), float(c[1] + h[1]), float(c[2]), m))
    w = s.list(objs)
    cam = rtb.camera((26, 6, 6), (0, 3, 0), (0, 1, 0), 30.0, nx / ny, 0.0, 10.0, 0.0, 1.0)
    return s, w, cam


def cornell_like(n_boxes, nx, ny):
    """The cornell_box room (tests/test.rs:242-323: five walls + light) with n rotated boxes inside: an ENCLOSED scene."""
    rng = np.random.default_rng(100 + n_boxes)
    s = rtb.Scene()
    red, white, green = (s.lambertian(s.solid(*c)) for c in ((0.65, 0.05, 0.05), (0.73, 0.73, 0.73), (0.12, 0.45, 0.15)))
    light = s.diffuse_light(s.solid(15, 15, 15))
    YZ, ZX, XY = 0, 1, 2
    objs = [s.flip_normals(s.rect(YZ, 0, 0, 555, 555, 555, green)), s.rect(YZ, 0, 0, 555, 555, 0, red),
            s.rect(ZX, 227, 213, 332, 343, 554, light), s.flip_normals(s.rect(ZX, 0, 0, 555, 555, 555, white)),
            s.rect(ZX, 0, 0, 555, 555, 0, white), s.flip_normals(s.rect(XY, 0, 0, 555, 555, 555, white))]
    for i in range(n_boxes):
        h = rng.uniform(30, 90, 3)
        c = rng.uniform((60, 0, 60), (400, 300, 400))
        objs.append(s.translate(s.rotate(1, s.cube((0, 0, 0), tuple(h * 2), white), float(rng.uniform(-30, 30))), tuple(c)))
    w = s.list(objs)
    cam = rtb.camera((278, 278, -800), (278, 278, 0), (0, 1, 0), 40.0, nx / ny, 0.0, 10.0, 0.0, 1.0)
    return s, w, cam


def three_way(s, w, cam, nx, ny, ns, bg):
    line = ""
    for width in (2, 4, 1):
        o = rtb.default_opts(background=bg)
        o.reserved[6] = width
        s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
        best = 1e9
        for _ in range(3):
            _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
            best = min(best, st["device_ms"])
        line += f" | {'flat' if width == 1 else 'bvh%d' % width} {best:6.2f} ms"
    return line


if len(sys.argv) > 1 and sys.argv[1] == "--cornell":
    for n in (0, 1, 2, 4, 6, 10, 14, 20, 26):
        s, w, cam = cornell_like(n, 512, 512)
        print(f"  room + {n:2d} boxes = {6 + n:2d} primitives" + three_way(s, w, cam, 512, 512, 16, 0), flush=True)
    sys.exit(0)

if len(sys.argv) > 1 and sys.argv[1] == "--sweep":
    nx, ny, ns = 960, 540, 16
    for kind in ("sphere", "cube", "rect"):
        for n in (2, 4, 6, 8, 12, 16, 24, 32, 48):
            s, w, cam = synthetic(n, kind, nx, ny)
            line = f"  {kind:6s} n={n:3d} "
            for width in (2, 4, 1):
                o = rtb.default_opts(background=1)
                o.reserved[6] = width
                s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
                best = 1e9
                for _ in range(3):
                    _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
                    best = min(best, st["device_ms"])
                line += f" | {'flat' if width == 1 else 'bvh%d' % width} {best:6.2f} ms"
            print(line, flush=True)
    sys.exit(0)

names = sys.argv[1:] or ["two_spheres", "two_perlin_sphere", "earth", "simple_light", "cornell_box", "cornell_smoke"]
for name in names:
    nx, ny, ns = CASES[name]
    s, w, cam = rtb.Scene.named(name, nx, ny)
    line = f"  {name:18s} {s.prim_count(w) if hasattr(s, 'prim_count') else '?':>5} prims "
    for width in (2, 4, 1):
        o = rtb.default_opts()
        o.reserved[6] = width
        s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
        best = 1e9
        for _ in range(3):
            _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
            best = min(best, st["device_ms"])
        line += f" | {'flat' if width == 1 else 'bvh%d' % width} {best:7.2f} ms {st['paths'] / best / 1e3:8.1f} Mpaths/s"
    print(line, flush=True)
