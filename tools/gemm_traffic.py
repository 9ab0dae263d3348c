#!/usr/bin/env python
"""profiles/gemm_traffic.json from an `ncu --set full` capture of the tcgen05 GEMM launches of one step:
   ncu -i rep --page raw --csv | python tools/gemm_traffic.py <source-name> > profiles/gemm_traffic.json"""
import csv, json, sys
rows = list(csv.reader(sys.stdin))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
def val(r, m):
    v = float(r[idx[m]].replace(",", ""))
    u = units[idx[m]].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
per = []
for r in data:
    if "gemm_bf16x2_kernel" not in r[idx["Kernel Name"]]:
        continue
    per.append({"kernel": r[idx["Kernel Name"]].split("(")[0].replace("void plr2::", ""),
                "us": float(r[idx["gpu__time_duration.sum"]]) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(units[idx["gpu__time_duration.sum"]], 1.0),
                "dram_read": val(r, "dram__bytes_read.sum"), "dram_write": val(r, "dram__bytes_write.sum")})
tot = sum(p["dram_read"] + p["dram_write"] for p in per)
print(json.dumps({"source": sys.argv[1] if len(sys.argv) > 1 else "", "launches": len(per),
                  "avg_dram_bytes_per_launch": tot / max(1, len(per)), "per_launch": per}, indent=1))
