#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals + sequence."""
import csv, re, sys
from collections import OrderedDict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1e3 if unit.startswith("ns") else (v if unit.startswith("us") else v * 1e3)
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = name.replace("plr2::", "")
    rows.append((name, us))
tot = sum(u for _, u in rows)
agg = OrderedDict()
for n, u in rows:
    a = agg.setdefault(n, [0.0, 0])
    a[0] += u; a[1] += 1
for n, (u, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{u:9.1f} us {100*u/tot:5.1f}%  x{c:3d}  {n}")
print(f"TOTAL {tot:.1f} us over {len(rows)} launches\n\nsequence:")
for n, u in rows:
    print(f"{u:8.1f}  {n}")
