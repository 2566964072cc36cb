"""Soak test: random image sizes / sample counts / shards / pipelines / thresholds on several scenes; every render must
finish (persistent kernels: no lost work item, no stuck WAIT state) and produce exactly nx*ny*samples paths.
usage: fuzz_sizes.py [seconds] [seed]"""
import sys, time, random
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import raytracing_rust_b200 as rtb

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rnd = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
names = ["cornell_box", "final_scene_lit", "random_scene", "cornell_smoke_book", "two_spheres", "simple_light"]
scenes = {}
t_end, n = time.time() + budget, 0
while time.time() < t_end:
    name = rnd.choice(names)
    nx, ny = rnd.choice([1, 2, 7, 8, 9, 31, 33, 64, 100, 257, 640]), rnd.choice([1, 3, 4, 5, 17, 32, 36, 99, 360])
    ns = rnd.choice([1, 2, 3, 5, 16, 63, 64, 65, 130])
    stride = rnd.choice([1, 1, 2, 3, 8]); begin = rnd.randrange(stride)
    pipe = rnd.choice([rtb.PIPE_MEGAKERNEL, rtb.PIPE_MEGAKERNEL, rtb.PIPE_WAVEFRONT])
    key = (name, nx, ny)
    s, w, cam = rtb.Scene.named(name, nx, ny)
    if rnd.random() < 0.3:
        s.set_bvh_builder(rtb.BVH_DEVICE_LBVH)
    o = rtb.default_opts(pipeline=pipe, background=rnd.choice([0, 1]), seed=rnd.randrange(1 << 30), sample_begin=begin, sample_stride=stride,
                         collect_stats=rnd.choice([0, 1]))
    rows = ny
    if pipe == rtb.PIPE_MEGAKERNEL:
        for i in range(3):
            o.reserved[i] = rnd.choice([0, 1, 4, 8, 16, 32])
        o.reserved[3] = rnd.choice([0, 2, 3, 4, 5, 6, 7])
        o.reserved[6] = rnd.choice([0, 1, 2, 4])  # tree: upload-time choice / flat scan / two-child / four-wide
        if rnd.random() < 0.4:  # a row-band shard on top of the sample shard
            o.band_stride = rnd.choice([2, 3, 8])
            o.band_begin = rnd.randrange(o.band_stride)
            rows = sum(min(4, ny - 4 * b) for b in range(o.band_begin, (ny + 3) // 4, o.band_stride))
    img, _, st = s.render(w, cam, nx, ny, ns, o, want_rgb8=False, want_stats=True)
    mine = len(range(begin, ns, stride))
    assert st["paths"] == nx * rows * mine, (name, nx, ny, ns, begin, stride, pipe, st["paths"])
    assert np.isfinite(img).all() and (img >= 0).all()
    n += 1
try:
    fails = rtb.check_failures()
    assert not any(fails.values()), fails
    print("self-check counters clean:", fails)
except rtb.RtbError:
    pass
print(f"fuzz ok: {n} renders in {budget:.0f} s")
