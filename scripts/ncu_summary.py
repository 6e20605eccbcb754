#!/usr/bin/env python
"""ncu --page raw --csv of a --set full capture -> a short per-kernel text summary (first captured launch of every
distinct kernel).  usage: ncu_summary.py raw.csv [header line ...] > profiles/rNN_ncu_full_*.txt"""
import csv
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]
rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
for line in sys.argv[2:]:
    print("# " + line)
seen = set()
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    key = name.split("(")[0]
    if key in seen:
        continue
    seen.add(key)
    print(name[:110])
    for m in METRICS:
        if m in ix:
            print(f"    {m:70s} {r[ix[m]]} {units[ix[m]]}")
