"""Hunts intermittent slow rtb_render calls (r1 verdict weak #6): N calls exactly as bench.py's e2e leg makes them (fresh scene
handle per call, host buffers) with RTB200_TIMING=1 stage laps; prints the stages of every call slower than 1.05x the median.
usage: e2e_stall_hunt.py [calls]"""
import os, statistics, subprocess, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
if os.environ.get("RTB_HUNT_CHILD"):
    sys.path.insert(0, str(ROOT))
    import raytracing_rust_b200 as rtb
    nx, ny, ns = 3840, 2160, 64
    s, w, cam = rtb.Scene.named("final_scene", nx, ny)
    blob = s.serialize()
    for k in range(int(sys.argv[1])):
        t0 = time.perf_counter()
        s2 = rtb.Scene.deserialize(blob)
        t1 = time.perf_counter()
        g, g8, _ = s2.render(w, cam, nx, ny, ns, rtb.default_opts(), want_sum=False)
        t2 = time.perf_counter()
        del s2
        t3 = time.perf_counter()
        print(f"[hunt] call {k}: total {1e3*(t3-t0):.1f} ms = deserialize {1e3*(t1-t0):.1f} + render {1e3*(t2-t1):.1f} + del {1e3*(t3-t2):.1f}",
              file=sys.stderr, flush=True)
    sys.exit(0)
n = sys.argv[1] if len(sys.argv) > 1 else "24"
r = subprocess.run([sys.executable, __file__, n], env=dict(os.environ, RTB_HUNT_CHILD="1", RTB200_TIMING="1"), capture_output=True, text=True)
calls, cur = [], []
for line in r.stderr.splitlines():
    cur.append(line)
    if line.startswith("[hunt] call"):
        calls.append((float(line.split("total ")[1].split(" ms")[0]), cur))
        cur = []
tot = [c[0] for c in calls]
med = statistics.median(tot)
print(f"{len(tot)} calls: median {med:.1f} ms, min {min(tot):.1f}, max {max(tot):.1f}, mean {statistics.mean(tot):.1f}; "
      f"calls slower than 1.05 x median: {sum(t > 1.05 * med for t in tot)}")
for t, lines in calls[1:]:
    if t > 1.05 * med:
        print(f"--- slow call ({t:.1f} ms)")
        for l in lines:
            print("   ", l)
print("--- a typical call")
for l in min(calls, key=lambda c: abs(c[0] - med))[1]:
    print("   ", l)
