import csv, sys, collections, re
f = sys.argv[1]
rows = list(csv.reader(open(f)))
# may contain several kernels: split on "Kernel Name" rows; take first
hdr = None; data = []
nk = 0
for r in rows:
    if r and r[0] == "Kernel Name":
        nk += 1
        if nk > 1: break
        continue
    if r and r[0] == "Address": hdr = r; continue
    if hdr and len(r) == len(hdr): data.append(r)
ix = {h:i for i,h in enumerate(hdr)}
tot = collections.Counter(); thr = collections.Counter(); samp = collections.Counter()
for r in data:
    s = r[ix["Source"]].strip()
    s = re.sub(r'^@!?U?P\d+\s+', '', s)
    op = s.split()[0].split('.')[0]
    n = int(r[ix["Instructions Executed"]]); t = int(r[ix["Thread Instructions Executed"]])
    tot[op] += n; thr[op] += t; samp[op] += int(r[ix["# Samples"]])
T = sum(tot.values()); S = sum(samp.values())
print("total warp inst", T, "avg thr", sum(thr.values())/T)
for op, n in tot.most_common(40):
    print(f"{op:10s} {n:12d} {100*n/T:6.2f}%  thr/inst {thr[op]/max(n,1):5.1f}  samples {100*samp[op]/S:5.1f}%")
