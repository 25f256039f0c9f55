#!/usr/bin/env python
"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel name: launches, total us, share.
usage: launch_agg.py launches.csv [first_id]   (first_id: ignore launches with ID below it, e.g. the warm-up pass)"""
import csv, sys, re, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
h = rows[0]; iid, ik, iv, iu = h.index("ID"), h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= iv or int(r[iid]) < first: continue
    v = float(r[iv].replace(",", "")); u = r[iu]
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
    name = re.sub(r"\(.*$", "", r[ik])
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':90s} {'n':>5s} {'total us':>10s} {'us/launch':>9s} {'share':>6s}")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:90]:90s} {n:5d} {t:10.1f} {t / n:9.1f} {t / tot:6.1%}")
print(f"{'total':90s} {sum(a[0] for a in agg.values()):5d} {tot:10.1f}")
