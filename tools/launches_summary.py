"""Condense an `ncu --metrics gpu__time_duration.sum --csv` launch list into kernel, launches, total ms, share."""
import csv, sys, collections
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    ms = v / 1e6 if r["Metric Unit"] in ("ns", "nsecond") else (v / 1e3 if r["Metric Unit"] in ("us", "usecond") else v)
    k = r["Kernel Name"][:80]
    tot[k][0] += 1
    tot[k][1] += ms
s = sum(v[1] for v in tot.values())
print("kernel,launches,total_ms,share")
for k, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f'"{k}",{n},{ms:.3f},{ms / s:.4f}')
