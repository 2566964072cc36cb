"""Join an ncu SASS source page (CSV) with nvdisasm line info and aggregate per CUDA source line.
usage: ncu_by_line.py <report.ncu-rep> <cubin> <kernel-mangled-substring> [top_n]"""
import csv, re, subprocess, sys, collections

rep, cubin, kname = sys.argv[1:4]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 50
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# collect (offset -> (file,line)) for the kernel section
in_k = False
cur = ("?", 0)
off2line = {}
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
hdr_i = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
hdr = rows[hdr_i]
base = None
agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter()])
tot_i = tot_t = tot_s = 0
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    a = int(d["Address"], 16)
    if base is None:
        base = a
    key = off2line.get(a - base, ("?", 0))
    ie, te, sm = int(d["Instructions Executed"]), int(d["Thread Instructions Executed"]), int(d["# Samples"])
    g = agg[key]
    g[0] += ie; g[1] += te; g[2] += sm
    for c in stall_cols:
        v = int(d[c] or 0)
        if v:
            g[3][c] += v
    tot_i += ie; tot_t += te; tot_s += sm
print(f"total warp-inst {tot_i:,}  thread-inst {tot_t:,}  avg active {tot_t/max(tot_i,1):.2f}  samples {tot_s}")
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), g in agg.items():
    byfile[f][0] += g[0]; byfile[f][1] += g[1]; byfile[f][2] += g[2]
for f, g in sorted(byfile.items(), key=lambda kv: -kv[1][2]):
    print(f"  {f:20s} inst {g[0]/tot_i*100:5.1f}%  samples {g[2]/tot_s*100:5.1f}%  active {g[1]/max(g[0],1):5.1f}")
src_cache = {}
def src(f, l):
    import glob
    if f not in src_cache:
        import os
        root = os.environ.get("RTB_SRC_ROOT", "/root/repo")  # the tree the profiled binary was built from
        c = glob.glob(f"{root}/**/{f}", recursive=True)
        src_cache[f] = open(c[0]).read().splitlines() if c else []
    L = src_cache[f]
    return L[l - 1].strip()[:90] if 0 < l <= len(L) else ""
for (f, l), g in sorted(agg.items(), key=lambda kv: -kv[1][2])[:topn]:
    st = ",".join(f"{k[6:]}:{v}" for k, v in g[3].most_common(3))
    print(f"{g[2]/tot_s*100:5.1f}% smp {g[0]/tot_i*100:5.1f}% inst act {g[1]/max(g[0],1):5.1f} {f}:{l:<4d} {src(f,l)}  [{st}]")
