import sys, time, os
sys.path.insert(0, '/root/repo')
import raytracing_rust_b200 as rtb
nx, ny, ns = 3840, 2160, 64
s, w, cam = rtb.Scene.named("final_scene", nx, ny)
s.render(w, cam, 64, 64, 1)
blob = s.serialize()
for rep in range(3):
    t0 = time.perf_counter(); s2 = rtb.Scene.deserialize(blob); t1 = time.perf_counter()
    g, g8, st = s2.render(w, cam, nx, ny, ns, rtb.default_opts(), want_sum=False); t3 = time.perf_counter()
    print(f"deserialize {1e3*(t1-t0):.2f} ms  render call {1e3*(t3-t1):.2f} ms total {1e3*(t3-t0):.1f}", file=sys.stderr, flush=True)
    t4=time.perf_counter(); del s2; print(f"del scene {1e3*(time.perf_counter()-t4):.2f} ms", file=sys.stderr)
