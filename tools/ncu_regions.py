"""Per-region table of one k_render_mega launch: the ncu SASS source page joined with nvdisasm line info and summed over
source FUNCTIONS (trace.cuh / philox.h) and the megakernel's phases (render.cu line ranges of the profiled build).
usage: RTB_SRC_ROOT=<tree the binary was built from> ncu_regions.py <report.ncu-rep> <cubin> <kernel-mangled-substring>"""
import bisect, collections, csv, os, re, subprocess, sys

rep, cubin, kname = sys.argv[1:4]
root = os.environ.get("RTB_SRC_ROOT", "/root/repo")
csrc = f"{root}/raytracing_rust_b200/csrc"
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
in_k, cur, off2line = False, ("?", 0), {}
for ln in dis:
    if ln.startswith("//---") and ".text." in ln:
        in_k = kname in ln
        continue
    if not in_k:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        off2line[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
hdr = rows[hi]


def func_ranges(path):
    rng = []
    for i, l in enumerate(open(path), 1):
        m = re.match(r"(?:RTB_DEV|__device__ __forceinline__|__global__|RTB_HD inline)\s+[\w:<>\*& ]+?\b(\w+)\(", l)
        if m:
            rng.append((i, m.group(1)))
    return rng


fr = {f: func_ranges(f"{csrc}/{f}") for f in ("trace.cuh", "philox.h")}
# phases of k_render_mega by the marker comments of render.cu
marks = []
for i, l in enumerate(open(f"{csrc}/render.cu"), 1):
    for key, name in (("int sample_media_const(", "sample_media_const"), ("__global__ void __launch_bounds__(kThreads", "kernel prologue"),
                      ("unsigned m_node = __ballot_sync", "NODE loop: vote / push / pop (render.cu)"), ("if (check_wait) {", "waiter test / item fetch"),
                      ("const unsigned m_leaf = __ballot_sync", "phase select"), ("SHADE phase", "SHADE body (render.cu)"),
                      ("lanes whose path ended take the next paths", "regeneration + segment start"), ("below-threshold NODE step", "below-threshold NODE step"),
                      ("LEAF phase", "LEAF loop (render.cu)"), ("  __syncwarp();\n", "epilogue (tile flush, counters)"), ("__global__ void __launch_bounds__(256) k_closest_hit", "other kernels")):
        if key in l and (not marks or marks[-1][1] != name):
            marks.append((i, name))


def bucket(f, l):
    if f == "render.cu":
        idx = bisect.bisect_right([x[0] for x in marks], l) - 1
        return marks[idx][1] if idx >= 0 else "render.cu (top)"
    if f in fr:
        r = fr[f]
        idx = bisect.bisect_right([x[0] for x in r], l) - 1
        return f + "::" + (r[idx][1] if idx >= 0 else "?")
    return f


agg, base, T = collections.defaultdict(lambda: [0, 0, 0]), None, [0, 0, 0]
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    a = int(d["Address"], 16)
    if base is None:
        base = a
    f, l = off2line.get(a - base, ("?", 0))
    ie, te, sm = int(d["Instructions Executed"]), int(d["Thread Instructions Executed"]), int(d["# Samples"])
    g = agg[bucket(f, l)]
    g[0] += ie; g[1] += te; g[2] += sm
    T[0] += ie; T[1] += te; T[2] += sm
print(f"total warp-inst {T[0]:,}  thread-inst {T[1]:,}  avg active lanes {T[1] / max(T[0], 1):.2f}  samples {T[2]:,}")
print(f"{'region':48s} {'time %':>7s} {'inst %':>7s} {'active lanes':>13s}")
for k, g in sorted(agg.items(), key=lambda kv: -kv[1][2]):
    if g[2] / T[2] < 0.002:
        continue
    print(f"{k:48s} {g[2] / T[2] * 100:7.1f} {g[0] / T[0] * 100:7.1f} {g[1] / max(g[0], 1):13.1f}")
