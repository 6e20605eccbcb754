#!/bin/bash
# iteration pass: GPU parity tests, then a short bench without the host-buffer / CPU arms
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_iter.json; tail -5 gpurun_out/bench_iter.err
