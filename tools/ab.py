"""A/B experiment builds of librtb200.so (see Makefile: B=variants/x LIB=variants/x/librtb200.so EXTRA=-D...).
usage: ab.py [--scenes a,b] [--stats] lib1.so lib2.so ...   (device-event ms, best of 3)"""
import os, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
CASES = {"final_scene": (1920, 1080, 16), "cornell_box": (1024, 1024, 32), "random_scene": (1920, 1080, 16),
         "two_perlin_sphere": (1920, 1080, 16), "cornell_smoke": (1024, 1024, 32), "final4k": (3840, 2160, 8),
         "two_spheres": (400, 225, 100), "simple_light": (1280, 720, 16), "earth": (1920, 1080, 16)}
if os.environ.get("RTB_AB_CHILD"):
    sys.path.insert(0, str(ROOT))
    import raytracing_rust_b200 as rtb
    scenes, stats = sys.argv[1].split(","), sys.argv[2] == "1"
    knobs = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 and sys.argv[3] else []
    for name in scenes:
        nx, ny, ns = CASES[name]
        s, w, cam = rtb.Scene.named("final_scene" if name == "final4k" else name, nx, ny)
        o = rtb.default_opts()
        for i, k in enumerate(knobs):
            o.reserved[i] = k
        s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
        best = 1e9
        for _ in range(3):
            _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
            best = min(best, st["device_ms"])
        line = f"  {name:18s} {best:8.2f} ms {st['paths']/best/1e3:8.1f} Mpaths/s"
        if stats:
            o2 = rtb.default_opts(collect_stats=1)
            _, _, s2 = s.render(w, cam, nx, ny, 4, o2, want_sum=False, want_rgb8=False, want_stats=True)
            seg = s2["segments"]
            line += f"  node/seg {s2['node_tests']/seg:.2f} prim/seg {s2['prim_tests']/seg:.2f} seg/path {seg/s2['paths']:.2f}"
        print(line, flush=True)
    sys.exit(0)
args = sys.argv[1:]
scenes, stats, knobs = "final_scene,cornell_box,random_scene", "0", ""
libs = []
while args:
    a = args.pop(0)
    if a == "--scenes": scenes = args.pop(0)
    elif a == "--stats": stats = "1"
    elif a == "--knobs": knobs = args.pop(0)
    else: libs.append(a)
for lib in libs:
    print(lib, flush=True)
    env = dict(os.environ, RTB_AB_CHILD="1", RTB200_LIB=str((ROOT / lib).resolve()))
    subprocess.run([sys.executable, __file__, scenes, stats, knobs], env=env)
