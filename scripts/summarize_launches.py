"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1], errors="ignore")) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
H = rows[hdr]
kn, mv, mu = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
tot = collections.defaultdict(lambda: [0.0, 0])
for r in rows[hdr + 1:]:
    try:
        v = float(r[mv].replace(",", ""))
    except ValueError:
        continue
    u = r[mu]
    ms = v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v
    name = r[kn].split("(")[0][:70]
    tot[name][0] += ms
    tot[name][1] += 1
total = sum(v[0] for v in tot.values())
print(f"total {total:.2f} ms, {sum(v[1] for v in tot.values())} launches")
for name, (ms, n) in sorted(tot.items(), key=lambda kv: -kv[1][0]):
    print(f"  {ms:8.2f} ms {100 * ms / total:5.1f}%  x{n:4d}  {name}")
