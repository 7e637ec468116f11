"""Sum an ncu --csv launch list (gpu__time_duration.sum) by kernel name: python scripts/summarize_launches.py file.csv [last_n]"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit")
body = rows[1:]
if len(sys.argv) > 2:
    body = body[-int(sys.argv[2]):]
tot = collections.Counter(); cnt = collections.Counter()
for r in body:
    v = float(r[vi].replace(",", "")) * (1e-3 if r[ui] in ("ns", "nsecond") else 1.0)
    tot[r[ki][:70]] += v; cnt[r[ki][:70]] += 1
s = sum(tot.values())
print(f"total {s:.1f} us, {len(body)} launches")
for k, v in tot.most_common(14):
    print(f"{v:10.1f} us {100 * v / s:5.1f}% x{cnt[k]:4d} {k}")
