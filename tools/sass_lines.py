"""Print the SASS of one kernel between two source lines of a file (with line info).
usage: sass_lines.py <lib.so> <kernel-substring> <file> <first_line> <last_line>"""
import re, subprocess, os, glob, sys, tempfile
obj, kname, f, lo, hi = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
out = []
for c in sorted(glob.glob(tmp + "/*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", c], capture_output=True, text=True).stdout.splitlines()
    in_k, cur = False, ("?", 0)
    for ln in dis:
        if ln.startswith("//---") and ".text." in ln:
            in_k = kname in ln
            continue
        if not in_k:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out.append((int(m.group(1), 16), cur, m.group(2)))
        elif re.match(r"\.L_x_\d+:", ln.strip()):
            out.append((-1, cur, ln.strip()))
idx = [i for i, (off, c, ins) in enumerate(out) if c[0] == f and lo <= c[1] <= hi]
if idx:
    for off, c, ins in out[min(idx):max(idx) + 1]:
        print(f"{off:6x} {c[0][:12]:12s}{c[1]:4d} {ins}")
print("total instructions in kernel:", sum(1 for o in out if o[0] >= 0))
