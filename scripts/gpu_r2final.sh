#!/bin/bash
# final pass of the round: whole GPU suite, smoke, then the measurement pass (scripts/gpu_r2prof.sh)
mkdir -p gpurun_out; rm -f gpurun_out/parity_r02.txt
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^FAILED|passed|failed" gpurun_out/pytest_gpu.log | tail -5
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
bash scripts/gpu_r2prof.sh
