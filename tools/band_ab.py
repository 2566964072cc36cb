"""One GPU's share of an 8-GPU 4K image, as a sample shard and as a row-band shard (same number of paths): device ms.
usage: band_ab.py [spp_total]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
nx, ny = 3840, 2160
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
s, w, cam = rtb.Scene.named("final_scene", nx, ny)
s.render(w, cam, nx, ny, 2, want_sum=False, want_rgb8=False)
for label, kw in (("sample shard 0/8", dict(sample_begin=0, sample_stride=8)), ("band shard 0/8", dict(band_begin=0, band_stride=8)),
                  ("band shard 3/8", dict(band_begin=3, band_stride=8))):
    best = 1e9
    for _ in range(2):
        _, _, st = s.render(w, cam, nx, ny, ns, rtb.default_opts(**kw), want_sum=False, want_rgb8=False, want_stats=True)
        best = min(best, st["device_ms"])
    print(f"{label:18s} {ns} spp image: {best:9.2f} ms  {st['paths'] / best / 1e3:8.1f} Mpaths/s  ({st['kernel_launches']} launches)", flush=True)
