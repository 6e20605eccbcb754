#!/bin/bash
mkdir -p gpurun_out
B200MP_TORCH_PROFILE=gpurun_out/tab_torch_profile.txt timeout 900 python bench.py --model tabgnn --steps 2 --warmup 2 --no-e2e --no-cpu-baseline > gpurun_out/tab_t.json 2> gpurun_out/tab_t.err; python scripts/benchsum.py gpurun_out/tab_t.json > gpurun_out/tab_t.txt; head -30 gpurun_out/tab_t.txt; tail -3 gpurun_out/tab_t.err
cut -c1-200 gpurun_out/tab_torch_profile.txt | head -50
