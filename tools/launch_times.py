"""Summarise an `ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --csv` launch list per kernel."""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith("==")]
h = rows[0]; kn = h.index("Kernel Name"); mn = h.index("Metric Name"); mv = h.index("Metric Value")
t = collections.OrderedDict()
for r in rows[1:]:
    d = t.setdefault(r[kn][:60], {"n": 0, "ns": 0.0, "inst": 0.0})
    v = float(r[mv].replace(",", ""))
    if r[mn] == "gpu__time_duration.sum": d["ns"] += v; d["n"] += 1
    elif r[mn] == "smsp__inst_executed.sum": d["inst"] += v
tot = sum(d["ns"] for d in t.values())
for k, d in t.items():
    print(f"{k:60s} x{d['n']:<3d} {d['ns']/1e6:9.3f} ms {100*d['ns']/tot:5.1f}%  {d['inst']:.3e} inst")
