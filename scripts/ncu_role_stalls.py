"""Warp-stall samples per WARP ROLE of a warp-specialised kernel, from `ncu -i X.ncu-rep --page source --csv
--print-source sass`: the roles are contiguous SASS regions (the role switch is one BRX / branch chain), so the rows
are cut at the given instruction indices.   usage: ncu_role_stalls.py sass.csv name:first_index [name:first_index ...]
(find the indices with the marker instructions: UBLKCP = producer, UTCHMMA = MMA issuers, MUFU.EX2 = softmax)"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
body = [r for r in rows[2:] if len(r) == len(hdr)]
cuts = [(a.split(":")[0], int(a.split(":")[1])) for a in sys.argv[2:]]
isamp, iex = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(int(r[isamp] or 0) for r in body)
print(f"{len(body)} SASS instructions, {tot} samples")
for k, (name, a) in enumerate(cuts):
    b = cuts[k + 1][1] if k + 1 < len(cuts) else len(body)
    seg = body[a:b]
    n = sum(int(r[isamp] or 0) for r in seg)
    ex = sum(int(r[iex] or 0) for r in seg)
    why = collections.Counter()
    for r in seg:
        for i in stall:
            why[hdr[i][6:]] += int(r[i] or 0)
    s = sum(why.values()) or 1
    mix = ", ".join(f"{w} {100 * c / s:.0f}%" for w, c in why.most_common(6) if c)
    print(f"{name:12s} SASS [{a:4d},{b:4d})  warp-instructions {ex:11d}  samples {n:6d} ({100 * n / tot:5.1f}%)  {mix}")
