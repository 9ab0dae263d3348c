#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into one row per launch with the metrics the roofline
discussion uses.  Usage: ncu -i rep --page raw --csv | python tools/ncu_summary.py"""
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "us"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor%"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed", "fma%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1%"),
        ("dram__bytes_read.sum", "dramR"), ("dram__bytes_write.sum", "dramW")]
cols = [(m, n) for m, n in cols if m in idx]
print(f"{'kernel':34s} " + " ".join(f"{n:>8s}" for _, n in cols))
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
NORM = {"ns": "us", "us": "us", "ms": "us", "s": "us", "byte": "Mbyte", "Kbyte": "Mbyte", "Mbyte": "Mbyte", "Gbyte": "Mbyte"}
print(f"{'':34s} " + " ".join(f"{NORM.get(units[idx[m]], units[idx[m]])[:8]:>8s}" for m, _ in cols))
for r in data:
    name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("plr2::", "")[:34]
    vals = []
    for m, _ in cols:
        try:
            vals.append(f"{float(r[idx[m]].replace(',', '')) * SCALE.get(units[idx[m]], 1.0):8.1f}")
        except ValueError:
            vals.append(f"{r[idx[m]]:>8s}")
    print(f"{name:34s} " + " ".join(vals))
