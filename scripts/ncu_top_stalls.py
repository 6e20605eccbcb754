#!/usr/bin/env python
"""Top stall sites from `ncu --page source --csv --print-source sass` output.
usage: ncu_top_stalls.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
# the file may hold several kernels; split on "Kernel Name" rows
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            body.append(rows[j]); j += 1
        si = hdr.index("# Samples"); so = hdr.index("Source")
        stall_cols = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[si] or 0) for r in body if len(r) > si)
        print(f"== {name[:90]}  total samples {tot}")
        agg = {}
        for k in stall_cols:
            agg[hdr[k]] = sum(int(r[k] or 0) for r in body if len(r) > k)
        print("   by reason:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
        for r in sorted(body, key=lambda r: -int(r[si] or 0))[:n]:
            why = sorted(((int(r[k] or 0), hdr[k]) for k in stall_cols), reverse=True)[:2]
            print(f"   {int(r[si] or 0):7d} {100.0 * int(r[si] or 0) / max(tot, 1):5.1f}%  {r[so][:70]:70s} {why}")
        i = j
    else:
        i += 1
