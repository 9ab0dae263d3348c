#!/usr/bin/env python
"""Top sampled SASS lines of one launch: ncu -i rep --page source --csv --kernel-name regex:X --launch-count 1 | ncu_hot.py [N]"""
import csv, sys
rows = list(csv.reader(sys.stdin))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
hdr = rows[1]
data = []
for r in rows[2:]:
    if len(r) > 10 and r[0].startswith("0x"):
        data.append(r)
    elif data and len(r) <= 2:
        break
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
tot = sum(int(r[i_s]) for r in data)
print(rows[0][1][:70], "| samples", tot, "| lines", len(data), "| warp-instr", sum(int(r[i_ex]) for r in data))
top = sorted(enumerate(data), key=lambda kv: -int(kv[1][i_s]))[:n]
for i, r in sorted(top):
    print(f"{i:5d} {int(r[i_s]):6d} {int(r[i_ex]):9d}  {r[i_src].strip()[:90]}")
