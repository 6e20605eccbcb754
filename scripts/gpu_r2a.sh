#!/bin/bash
# round 2, first pass: GPU tests (incl. the scale parity tests), bench line, occupancy diagnostics
mkdir -p gpurun_out; rm -f gpurun_out/parity_r02.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
B200MP_DEBUG_OCC=1 timeout 600 python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/occ.json 2> gpurun_out/occ.err; echo "occ rc=$?"
sort gpurun_out/occ.err | uniq -c | sort -rn | head -20
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python scripts/benchsum.py gpurun_out/bench.json 2>/dev/null | head -40
