#!/bin/bash
# Full measurement pass: parity tests, bench line (with e2e + CPU baseline), reference arm, ncu launch
# list of the bench command, ncu --set full capture of the per-edge / node kernels on hi_small.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
CMD2="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small"
$CMD2 > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'umma|seg_light' -s ${SKIP:-108} -c ${COUNT:-36} -f -o gpurun_out/prof_small $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full.log
# reports are too big to travel back (64 MiB cap): export the pages here, drop the .ncu-rep
ncu -i gpurun_out/prof_small.ncu-rep --page raw --csv > gpurun_out/prof_small_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_small.ncu-rep --page source --csv --print-source sass > gpurun_out/prof_small_src.csv 2>/dev/null
rm -f gpurun_out/prof_small.ncu-rep
# DRAM traffic of the per-edge kernels at the bench workload itself (one timed step: launches after the 3 warm-ups)
timeout 900 ncu --set full --clock-control none -k regex:'umma|seg_light' -s ${TSKIP:-108} -c 36 -f -o gpurun_out/prof_medium $CMD > gpurun_out/ncu_medium.log 2>&1
echo "ncu medium rc=$?"; tail -2 gpurun_out/ncu_medium.log
ncu -i gpurun_out/prof_medium.ncu-rep --page raw --csv > gpurun_out/prof_medium_raw.csv 2>/dev/null
rm -f gpurun_out/prof_medium.ncu-rep
# tabular-side kernels (rows a5, a12): timings at 4 M rows, ncu --set full at 0.5 M rows
timeout 300 python scripts/kbench_tab.py > gpurun_out/kbench_tab.log 2>&1; tail -9 gpurun_out/kbench_tab.log | head -8
timeout 600 ncu --set full --clock-control none -k regex:'col_' -c 12 -f -o gpurun_out/prof_tab python scripts/kbench_tab.py --rows 500000 --reps 1 > gpurun_out/ncu_tab.log 2>&1
echo "ncu tab rc=$?"
ncu -i gpurun_out/prof_tab.ncu-rep --page raw --csv > gpurun_out/prof_tab_raw.csv 2>/dev/null
rm -f gpurun_out/prof_tab.ncu-rep
du -sh gpurun_out; ls -la gpurun_out
