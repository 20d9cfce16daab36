"""Per-kernel table from an `ncu --metrics ... --csv --log-file` launch list: launches, mean/total duration, share of the
step, tensor-pipe activity and DRAM bytes.  usage: python tools/launch_table.py launches.csv [steps]"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if r]
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hi]
ix = {k: h.index(k) for k in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value")}
per = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= ix["Metric Value"]:
        continue
    key = r[ix["ID"]]
    d = per.setdefault(key, {"name": r[ix["Kernel Name"]]})
    try:
        v = float(r[ix["Metric Value"]].replace(",", ""))
    except ValueError:
        continue
    unit = r[ix["Metric Unit"]]
    m = r[ix["Metric Name"]]
    if m.startswith("gpu__time_duration"):
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
    if m.startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
    d[m] = v
agg = collections.OrderedDict()
for d in per.values():
    n = d["name"].split("(")[0].replace("void ", "").replace("flowcuda::", "")
    a = agg.setdefault(n, {"n": 0, "us": 0.0, "tc": 0.0, "dram": 0.0})
    a["n"] += 1
    a["us"] += d.get("gpu__time_duration.sum", 0.0)
    a["tc"] += d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 0.0) * d.get("gpu__time_duration.sum", 0.0)
    a["dram"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
tot = sum(a["us"] for a in agg.values())
steps = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
print(f"{'kernel':58s} {'launches':>8s} {'mean us':>9s} {'us/step':>9s} {'share':>6s} {'tensor%':>8s} {'DRAM MB/step':>13s}")
for n, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
    print(f"{n[:58]:58s} {a['n']:8d} {a['us'] / a['n']:9.2f} {a['us'] / steps:9.1f} {100 * a['us'] / tot:5.1f}% {a['tc'] / max(a['us'], 1e-9):8.1f} {a['dram'] / steps / 1e6:13.1f}")
print(f"{'total':58s} {sum(a['n'] for a in agg.values()):8d} {'':9s} {tot / steps:9.1f} {'':6s} {'':8s} {sum(a['dram'] for a in agg.values()) / steps / 1e6:13.1f}")
