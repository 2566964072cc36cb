import csv, subprocess, sys, collections, re
rep=sys.argv[1]
out = subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hi=next(i for i,r in enumerate(rows) if "Address" in r and "Source" in r)
hdr=rows[hi]
c=collections.Counter(); t=collections.Counter(); tot=0
for r in rows[hi+1:]:
    if len(r)<len(hdr): continue
    d=dict(zip(hdr,r))
    s=d["Source"].strip()
    s=re.sub(r"^@!?U?P\d+\s+","",s)
    op=s.split()[0].split(".")[0] if s else "?"
    ie=int(d["Instructions Executed"]); te=int(d["Thread Instructions Executed"])
    c[op]+=ie; t[op]+=te; tot+=ie
for op,n in c.most_common(40):
    print(f"{op:12s} {n/tot*100:6.2f}%  active {t[op]/max(n,1):5.1f}")
