#!/bin/bash
# round 2, pass d: full GPU test suite with the fixes, occupancy limiter from ncu, forced 2 CTAs/SM experiment, bench
mkdir -p gpurun_out; rm -f gpurun_out/parity_r02.txt
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^FAILED|passed|failed" gpurun_out/pytest_gpu.log | tail -12
grep -E "config 1 \(" gpurun_out/pytest_gpu.log | head -2
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small"
timeout 600 ncu --section Occupancy --section LaunchStats --clock-control none -k regex:'edge_linear_fwd_umma_kernel' -c 1 $CMD > gpurun_out/ncu_occ.txt 2>&1
grep -E "Limit|Occupancy|Registers|Shared Memory|Block Size|Grid Size|Waves" gpurun_out/ncu_occ.txt | head -30
for f in 1 2; do
  B200MP_FORCE_CTAS_PER_SM=$f timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/force_$f.json 2>/dev/null
  echo "force ctas/sm=$f"; python scripts/benchsum.py gpurun_out/force_$f.json | grep -E "ms/step|message_fwd|node_linear64|node_linear "
done
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python scripts/benchsum.py gpurun_out/bench.json 2>/dev/null | head -40
