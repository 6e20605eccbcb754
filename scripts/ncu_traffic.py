#!/usr/bin/env python
"""ncu --page raw --csv of a --set full capture -> profiles/traffic.json (DRAM bytes per launch per
kernel, averaged over the captured launches).  usage: ncu_traffic.py raw.csv workload out.json"""
import csv, json, sys
NAMES = {"edge_update_bwd_umma_kernel": "edge_update_bwd", "message_bwd_umma_kernel<0": "pna_message_bwd",
         "message_bwd_umma_kernel<1": "gine_message_bwd", "message_bwd_umma64_kernel<0": "pna_message_bwd",
         "message_bwd_umma64_kernel<1": "gine_message_bwd", "edge_linear_fwd_umma_kernel<0": "pna_message_fwd",
         "edge_linear_fwd_umma_kernel<1": "gine_message_fwd", "edge_linear_fwd_umma_kernel<2": "edge_update_fwd",
         "seg_light_kernel<1": "segment_reduce_pna"}
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
def to_bytes(v, u):
    return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]
acc = {}
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    key = next((v for k, v in NAMES.items() if k in name), None)
    if key is None:
        continue
    b = sum(to_bytes(r[ix[m]], units[ix[m]]) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    ms = float(r[ix["gpu__time_duration.sum"]]) * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}[units[ix["gpu__time_duration.sum"]]]
    acc.setdefault(key, []).append((b, ms))
out = {"workload": sys.argv[2], "source": "ncu --set full --clock-control none, bench.py --steps 1 --warmup 3 (cold-cache replays)",
       "kernels": {k: {"dram_bytes_per_launch": round(sum(b for b, _ in v) / len(v)), "launches": len(v),
                       "ms_under_ncu": round(sum(m for _, m in v) / len(v), 3)} for k, v in acc.items()}}
json.dump(out, open(sys.argv[3], "w"), indent=1)
print(json.dumps(out))
