#!/bin/bash
# parity tests, then the kernel micro-benchmarks (scripts/kbench.py)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 500 python scripts/kbench.py ${KB_ARGS} > gpurun_out/kbench.log 2>&1; grep " ms" gpurun_out/kbench.log
