"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (second half = the last call)."""
import collections, csv, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
rows = rows[1:]
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
rows = rows[int(len(rows) * (1 - frac)):]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    if u in ("ns", "nsecond"):
        v /= 1e3
    elif u in ("ms", "msecond"):
        v *= 1e3
    k = r[ki][:70]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1]:10.1f} us {v[0]:5d} {100 * v[1] / tot:5.1f}%  {k}")
print(f"{tot:10.1f} us total over {len(rows)} launches")
