#!/bin/bash
# round 2 measurement pass: bench line (e2e + CPU baseline + whole model), reference arm, ncu launch list of the bench command,
# ncu --set full of the per-edge / node kernels on hi_small and (DRAM traffic) on hi_medium, SASS instruction summary
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python scripts/benchsum.py gpurun_out/bench.json | head -24
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --model gine --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_gine.json 2>/dev/null; python scripts/benchsum.py gpurun_out/bench_gine.json | head -2
timeout 600 python bench.py --workload hi_small --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_small.json 2>/dev/null; python scripts/benchsum.py gpurun_out/bench_small.json | head -2
timeout 600 python bench.py --dtype bf16 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_bf16.json 2>/dev/null; python scripts/benchsum.py gpurun_out/bench_bf16.json | head -2
for f in 1 0; do   # H = 128 (pna.py:50 default width): tcgen05 composition (1) against the fused SIMT kernels (0)
  B200MP_TC128=$f timeout 600 python bench.py --workload hi_small --hidden 128 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_small_h128_tc$f.json 2>/dev/null; python scripts/benchsum.py gpurun_out/bench_small_h128_tc$f.json | head -1
done
timeout 900 python bench.py --model tabgnn --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tabgnn_small.json 2>/dev/null; python scripts/benchsum.py gpurun_out/bench_tabgnn_small.json | head -1
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
CMD2="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small"
$CMD2 > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'umma|seg_light' -s ${SKIP:-108} -c ${COUNT:-36} -f -o gpurun_out/prof_small $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full.log
ncu -i gpurun_out/prof_small.ncu-rep --page raw --csv > gpurun_out/prof_small_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_small.ncu-rep --page source --csv --print-source sass > gpurun_out/prof_small_src.csv 2>/dev/null
rm -f gpurun_out/prof_small.ncu-rep
timeout 900 ncu --set full --clock-control none -k regex:'umma|seg_light' -s ${TSKIP:-108} -c 36 -f -o gpurun_out/prof_medium $CMD > gpurun_out/ncu_medium.log 2>&1
echo "ncu medium rc=$?"; tail -2 gpurun_out/ncu_medium.log
ncu -i gpurun_out/prof_medium.ncu-rep --page raw --csv > gpurun_out/prof_medium_raw.csv 2>/dev/null
rm -f gpurun_out/prof_medium.ncu-rep
du -sh gpurun_out
