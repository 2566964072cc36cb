"""k_render_dual against k_render_mega: same images (up to float add order), same path statistics; then a knob sweep.
usage: dual_check.py [knobs "k_node,k_leaf,k_shade,blocks[,_,carveout]" ...]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import raytracing_rust_b200 as rtb

for name, nx, ny, ns in (("cornell_box", 256, 256, 16), ("final_scene_lit", 320, 180, 16), ("two_spheres", 200, 112, 8),
                         ("cornell_smoke", 200, 200, 8), ("random_scene", 203, 117, 5)):
    s, w, cam = rtb.Scene.named(name, nx, ny)
    bg = rtb.BG_SKY if name in ("two_spheres", "random_scene") else rtb.BG_BLACK
    a, _, sa = s.render(w, cam, nx, ny, ns, rtb.default_opts(pipeline=rtb.PIPE_MEGAKERNEL, background=bg, collect_stats=1), want_stats=True)
    b, _, sb = s.render(w, cam, nx, ny, ns, rtb.default_opts(pipeline=rtb.PIPE_DUAL, background=bg, collect_stats=1), want_stats=True)
    print(name, "mean mega", a.mean() / ns, "dual", b.mean() / ns, "max abs diff", np.abs(a - b).max(), "segments",
          sa["segments"], sb["segments"], "node", sa["node_tests"], sb["node_tests"], flush=True)

nx, ny, ns = 1920, 1080, 16
s, w, cam = rtb.Scene.named("final_scene", nx, ny)
def run(pipe, knobs):
    o = rtb.default_opts(pipeline=pipe)
    for i, k in enumerate(knobs):
        o.reserved[i] = k
    s.render(w, cam, nx, ny, 2, o, want_sum=False, want_rgb8=False)
    best = 1e9
    for _ in range(3):
        _, _, st = s.render(w, cam, nx, ny, ns, o, want_sum=False, want_rgb8=False, want_stats=True)
        best = min(best, st["device_ms"])
    return best, st["paths"] / best / 1e3
ms, r = run(rtb.PIPE_MEGAKERNEL, ())
print(f"mega default: {ms:8.2f} ms {r:8.1f} Mpaths/s ", flush=True)
for knobs in [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]] or [(12, 12, 12, 2)]:
    ms, r = run(rtb.PIPE_DUAL, knobs)
    print(f"dual knobs {knobs}: {ms:8.2f} ms {r:8.1f} Mpaths/s", flush=True)
