"""Tiny runs of EVERY kernel of librtb200.so for compute-sanitizer (memcheck / racecheck / initcheck / synccheck):
the render pipelines (megakernel on the two-child and the four-wide tree, wavefront), row-band and sample
shards, the closest-hit / field-level hook kernels, the resolve / reduce-resolve / band kernels and the device LBVH builder.
usage: compute-sanitizer --tool <tool> python tools/sanitize_target.py [pipes]   (pipes: comma list, default 1,2)
   or: RTB200_LIB=exp/checks/librtb200.so python tools/sanitize_target.py   (the -DRTB_DEBUG_CHECKS=1 build: prints the
       kernels' own bounds-check counters — the stand-in where compute-sanitizer is not offered)"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import raytracing_rust_b200 as rtb

pipes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 2]
for name, nx, ny, ns in (("cornell_smoke_book", 37, 23, 6), ("final_scene_lit", 40, 24, 4), ("random_scene", 33, 17, 4)):
    s, w, cam = rtb.Scene.named(name, nx, ny)
    for p in pipes:
        a, r8, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(pipeline=p, background=1), want_stats=True)
        print(name, "pipe", p, "mean", float(a.mean()) / ns, "segments", st["segments"], flush=True)
    o = rtb.default_opts(background=1, collect_stats=1)
    for width in (2, 4, 1):  # the megakernel on either tree and on the flat scan, whatever the upload-time choice was
        o.reserved[6] = width
        a, _, st = s.render(w, cam, nx, ny, ns, o, want_stats=True)
        print(name, "mega/flat" if width == 1 else f"mega/bvh{width}", float(a.mean()) / ns, st["segments"], flush=True)
    a, _, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(background=1, band_begin=1, band_stride=3, sample_begin=1, sample_stride=2), want_stats=True)
    print(name, "band+sample shard", float(a.mean()) / ns, st["paths"], flush=True)
    rays = rtb.primary_rays(cam, nx, ny)
    prim, rec = s.closest_hit_ex(w, rays, self_id=None)
    prim2, _ = s.closest_hit_ex(w, rays, self_id=np.maximum(prim, -1))
    print(name, "closest hit", int((s.closest_hit(w, rays)[0] >= 0).sum()), int((prim >= 0).sum()), int((prim2 >= 0).sum()), flush=True)
# the no-allocation query (static device scratch)
import torch
d_rays = torch.from_numpy(rays.astype(np.float32)).cuda()
d_self = torch.from_numpy(np.maximum(prim, -1).astype(np.int32)).cuda()
d_prim = torch.empty(len(rays), dtype=torch.int32, device="cuda")
d_rec = torch.empty(len(rays), 9, dtype=torch.float32, device="cuda")
for _ in range(2):
    s.closest_hit_noalloc(d_rays.data_ptr(), len(rays), d_prim.data_ptr(), d_rec.data_ptr(), d_self.data_ptr())
torch.cuda.synchronize()
print("closest_hit_noalloc", int((d_prim >= 0).sum()))
# field-level hooks
print("camera", float(rtb.device_camera_rays(cam, 16, 9).sum()), "samplers", float(rtb.device_sample(0, 1, 1000).sum()), float(rtb.device_sample(1, 1, 1000).sum()))
s = rtb.Scene()
tex = s.checker(s.solid(0.2, 0.3, 0.1), s.noise(4.0, seed=7))
mats = [s.lambertian(tex), s.metal(s.solid(0.8, 0.6, 0.2), 0.3), s.dielectric(1.5), s.isotropic(s.image(rtb.earth_rgb8()[::16, ::16].copy())),
        s.diffuse_light(s.solid(7, 7, 7))]
w = s.list([s.sphere((0, 0, 0), 1.0, m) for m in mats])
inp = np.tile(np.array([0.7, 3.0, 2.8, 0.3, -1.0, 0.2, 0.25, 1.0, 0.3, 0.6, 1.0, 2.0, 3.0, 0.0, 1.0, 0.0]), (500, 1))
for m in mats:
    print("scatter", m, float(s.scatter(w, m, inp).sum()))
print("texture", float(s.texture_value(w, tex, np.random.default_rng(0).uniform(-5, 5, (500, 5))).sum()))
# exchange kernels on local buffers
import torch
nx, ny, ns = 37, 19, 8
bufs = [rtb.DeviceBuffer((ny, nx, 3)) for _ in range(3)]
for b in bufs:
    torch.as_tensor(b, device="cuda").fill_(2.0)
rgb8 = torch.zeros(ny, nx, 3, dtype=torch.uint8, device="cuda")
tot = torch.zeros(ny, nx, 3, dtype=torch.float32, device="cuda")
rtb.reduce_resolve_device([b.ptr for b in bufs], nx, ny, ns, rgb8.data_ptr(), tot.data_ptr())
n = 3 * nx * ny
rtb.reduce_resolve_range_device([b.ptr for b in bufs], 1000, n - 1000, ns, rgb8.data_ptr(), tot.data_ptr())
rtb.reduce_resolve_range_device([b.ptr for b in bufs], 0, 1000, ns, rgb8.data_ptr(), 0)
rtb.resolve_bands_device(bufs[0].ptr, nx, ny, ns, 1, 2, rgb8.data_ptr(), compact=False)
packed = torch.zeros(n, dtype=torch.uint8, device="cuda")
rtb.resolve_bands_device(bufs[0].ptr, nx, ny, ns, 0, 2, packed.data_ptr(), compact=True)
rtb.resolve_device(bufs[0].ptr, nx, ny, ns, rgb8.data_ptr())
torch.cuda.synchronize()
print("exchange kernels", int(rgb8.sum()), float(tot.sum()))
# device LBVH builder
s, w, cam = rtb.Scene.named("random_scene", 24, 16)
s.set_bvh_builder(rtb.BVH_DEVICE_LBVH)
a, _, st = s.render(w, cam, 24, 16, 4, rtb.default_opts(background=1), want_stats=True)
print("lbvh", float(a.mean()), st["segments"])
try:
    fails = rtb.check_failures()
    print("self-check counters (RTB_DEBUG_CHECKS build):", fails)
    assert not any(fails.values()), fails
    print("self-checks clean")
except rtb.RtbError as e:
    print("release build, no device self-checks:", e)
print("sanitize target done")
