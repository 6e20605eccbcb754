#!/bin/bash
# ncu full capture of our kernels on the hi_small workload (one timed step after 3 warm-ups)
mkdir -p gpurun_out
CMD2="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small"
$CMD2 > gpurun_out/plain2.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'edge_|seg_|message_' -s ${SKIP:-36} -c ${COUNT:-12} -f -o gpurun_out/${OUT:-prof_umma} $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
