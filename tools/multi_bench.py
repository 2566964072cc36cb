"""rtb_render_multi (one process, one host call) on 1..G devices: host wall clock of the whole call, host buffers.
usage: multi_bench.py [nx ny ns [only_this_device_count [reps]]]"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
nx, ny, ns = (int(x) for x in sys.argv[1:4]) if len(sys.argv) > 3 else (3840, 2160, 256)
s, w, cam = rtb.Scene.named("final_scene", nx, ny)
G = rtb.device_count()
only = int(sys.argv[4]) if len(sys.argv) > 4 else 0
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
s.render_multi(w, cam, 256, 144, G, devices=G, want_sum=False)  # contexts, replicas, uploads
g = only or 1
while g <= G:
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        _, img, st = s.render_multi(w, cam, nx, ny, ns, devices=g, want_sum=False, want_stats=True)
        best = min(best, time.perf_counter() - t0)
    print(f"devices {g}: {best*1e3:9.1f} ms per {nx}x{ny}x{ns} image  {nx*ny*ns/best/1e6:9.1f} Mpaths/s end to end "
          f"(slowest shard {st['device_ms']:.1f} ms on the device, {st['kernel_launches']} launches)", flush=True)
    if only:
        break
    g *= 2
