"""Print the SASS source page of one launch of an .ncu-rep: per-instruction executed counts and stall samples.
usage: ncu_src.py <rep> <kernel regex> <launch index> [min executed]"""
import csv, subprocess, sys, io
rep, rx, idx = sys.argv[1], sys.argv[2], int(sys.argv[3])
mn = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", f"regex:{rx}",
                      "--launch-skip", str(idx), "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
his = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hi = his[0]
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
tot = 0
recs = []
for r in rows[hi + 1:(his[1] if len(his) > 1 else len(rows))]:
    if len(r) < len(hdr):
        continue
    try:
        ie = int(r[ix["Instructions Executed"]]); sa = int(r[ix["Warp Stall Sampling (All Samples)"]])
        te = int(r[ix["Thread Instructions Executed"]])
    except ValueError:
        continue
    tot += sa
    recs.append((r[ix["Source"]].strip(), ie, te, sa))
print(rows[0][:2] if rows else "", "total samples", tot)
for s, ie, te, sa in recs:
    if ie >= mn:
        print(f"{s[:60]:60s} {ie:9d} {te / max(ie, 1):5.1f} {sa:5d}")
