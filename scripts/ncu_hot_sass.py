"""Hot SASS instructions of one kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass` output.
usage: ncu_hot_sass.py file.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.6
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
data = []
for r in rows[h + 1:]:
    if len(r) != len(hdr) or r[0] == "Address":
        if r and r[0] == "Kernel Name":
            break
        continue
    data.append(r)
isrc, isamp = hdr.index("Source"), hdr.index("# Samples")
stall_cols = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(int(r[isamp] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
for k, r in enumerate(data):
    n = int(r[isamp] or 0)
    if n > tot * min_pct / 100:
        top = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
        why = ", ".join(f"{w}:{c}" for c, w in top if c)
        print(f"{k:5d} {n:7d} {100 * n / tot:5.1f}%  {r[isrc][:90]:90s} {why}")
