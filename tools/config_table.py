"""BASELINE.json's five configs rendered in FULL on one B200 (time-to-image measured, not extrapolated), next to the CPU
oracle timed on a bounded sample of the same scene/camera on all host threads.  Writes a markdown table to stdout."""
import os, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import raytracing_rust_b200 as rtb
from oracle import oracle_py as orc

CONFIGS = [("C1", "two_spheres", 400, 225, 100), ("C2", "random_scene", 1920, 1080, 500), ("C3", "two_perlin_sphere", 1920, 1080, 1000),
           ("C3", "earth", 1920, 1080, 1000), ("C4", "cornell_box", 1024, 1024, 10000), ("C4", "cornell_smoke", 1024, 1024, 10000),
           ("C5", "final_scene", 3840, 2160, 10000)]
cpu_s = float(sys.argv[1]) if len(sys.argv) > 1 else 8.0
nth = os.cpu_count() or 1
print(f"| cfg | scene (as written, black background) | image | spp | paths | GPU s | GPU Mpaths/s | GPU Mseg/s | seg/path | CPU oracle Mpaths/s ({nth} thr) | CPU sample | speed-up |")
print("|---|---|---|---:|---:|---:|---:|---:|---:|---:|---|---:|")
for cfg, name, nx, ny, ns in CONFIGS:
    s, w, cam = rtb.Scene.named(name, nx, ny)
    s.render(w, cam, nx, ny, 2, want_sum=False, want_rgb8=False)
    t0 = time.perf_counter()
    _, rgb8, st = s.render(w, cam, nx, ny, ns, want_sum=False, want_stats=True)
    wall = time.perf_counter() - t0
    paths = nx * ny * ns
    # CPU oracle: same scene + camera at 1/16 of the pixels, spp sized for ~cpu_s seconds
    cx, cy = max(16, nx // 4), max(16, ny // 4)
    s2, w2, cam2 = rtb.Scene.named(name, cx, cy)
    o = orc.OracleScene(s2.serialize(), w2)
    t1 = time.perf_counter(); o.render(cam2, cx, cy, 1, seed=5, nthreads=nth); per = max(time.perf_counter() - t1, 1e-3)
    spp = max(1, min(2048, int(cpu_s / per)))
    t1 = time.perf_counter(); o.render(cam2, cx, cy, spp, seed=6, nthreads=nth); dt = time.perf_counter() - t1
    cpu = cx * cy * spp / dt / 1e6
    gpu = paths / (st["device_ms"] * 1e-3) / 1e6
    print(f"| {cfg} | {name} | {nx}x{ny} | {ns} | {paths/1e9:.3f} G | {st['device_ms']/1e3:.2f} (wall {wall:.2f}) | {gpu:.0f} | "
          f"{st['segments']/st['device_ms']/1e3:.0f} | {st['segments']/st['paths']:.2f} | {cpu:.2f} | {cx}x{cy}x{spp} in {dt:.1f} s | {gpu/cpu:.0f}x |", flush=True)
