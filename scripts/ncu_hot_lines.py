"""Stall samples per CUDA source line from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`.
usage: ncu_hot_lines.py file.csv [min_pct]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
cur_file, hdr = None, None
agg = collections.OrderedDict()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        isamp = hdr.index("# Samples")
        stall = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
        continue
    if hdr is None or len(r) != len(hdr) or not r[0]:
        continue  # SASS rows (empty line number) are already summed into their source line row
    key = (cur_file, int(r[0]))
    n = int(r[isamp] or 0)
    top = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall), reverse=True)[:3]
    e = agg.setdefault(key, [0, r[1].strip(), collections.Counter()])
    e[0] += n
    for c, w in top:
        e[2][w] += c
tot = sum(e[0] for e in agg.values())
print("total samples", tot)
for (f, ln), (n, src, why) in agg.items():
    if n >= tot * min_pct / 100:
        w = ", ".join(f"{k}:{v}" for k, v in why.most_common(2) if v)
        print(f"{f}:{ln:<4d} {n:6d} {100 * n / tot:5.1f}%  {src[:80]:80s} {w}")
