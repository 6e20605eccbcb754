#!/bin/bash
# round 2 debugging pass: precision breakdown, occupancy attributes, A/B timings of the new kernels on hi_small
mkdir -p gpurun_out
timeout 900 python scripts/debug_r2.py > gpurun_out/debug_r2.log 2>&1; echo "debug rc=$?"; grep -v Warning gpurun_out/debug_r2.log | tail -14
B200MP_DEBUG_OCC=1 timeout 300 python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > /dev/null 2> gpurun_out/occ2.err
sort gpurun_out/occ2.err | uniq -c | sort -rn | head -8
for cfg in "1 1" "0 1" "1 0" "0 0"; do set -- $cfg
  B200MP_FUSED_REDUCE=$1 B200MP_UPD_PIPE=$2 timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/ab_$1$2.json 2>/dev/null
  echo "fused=$1 pipe=$2"; python scripts/benchsum.py gpurun_out/ab_$1$2.json | grep -E "ms/step|message_fwd|segment_reduce|edge_update_fwd"
done
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'message_fwd_reduce|tile_records|zero_empty|edge_update_fwd_pipe' -c 24 --csv --log-file gpurun_out/launches_new.csv $CMD > /dev/null 2>&1
grep -E "message_fwd_reduce|tile_records|zero_empty|fwd_pipe" gpurun_out/launches_new.csv | awk -F'","' '{print $5, $NF}' | sort | uniq -c | head -20
