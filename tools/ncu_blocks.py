import csv, sys, re
rows=list(csv.reader(open(sys.argv[1])))
hdr=None; data=[]; nk=0
for r in rows:
    if r and r[0]=="Kernel Name":
        nk+=1
        if nk>1: break
        continue
    if r and r[0]=="Address": hdr=r; continue
    if hdr and len(r)==len(hdr): data.append(r)
ix={h:i for i,h in enumerate(hdr)}
T=sum(int(r[ix["Instructions Executed"]]) for r in data)
S=sum(int(r[ix["# Samples"]]) for r in data)
# group
groups=[]
cur=None
for i,r in enumerate(data):
    n=int(r[ix["Instructions Executed"]]); s=int(r[ix["# Samples"]]); t=int(r[ix["Thread Instructions Executed"]])
    if cur and cur["n"]==n:
        cur["cnt"]+=1; cur["s"]+=s; cur["t"]+=t; cur["ops"].append(r[ix["Source"]].strip().split()[0] if not r[ix["Source"]].strip().startswith('@') else r[ix["Source"]].strip().split()[1])
    else:
        cur={"start":i,"n":n,"cnt":1,"s":s,"t":t,"ops":[r[ix["Source"]].strip().split()[0]]}
        groups.append(cur)
print("total",T,"samples",S)
for g in groups:
    share=g["n"]*g["cnt"]/T
    if share>0.003 or g["s"]/S>0.005:
        import collections
        c=collections.Counter(o.split('.')[0] for o in g["ops"])
        print(f"@{g['start']:5d} x{g['cnt']:4d} exec {g['n']:9d} share {100*share:5.2f}% samp {100*g['s']/S:5.2f}% thr {g['t']/max(1,g['n']*g['cnt']):4.1f} ", dict(c.most_common(8)))
