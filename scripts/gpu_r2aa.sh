#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_col_attention.py tests/test_column_stack.py tests/test_gpu_models.py -m gpu -q > gpurun_out/pytest_aa.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_aa.log | tail -8
timeout 900 python bench.py --model tabgnn --steps 3 --warmup 2 --no-e2e --no-cpu-baseline > gpurun_out/tab_aa.json 2> gpurun_out/tab_aa.err; python scripts/benchsum.py gpurun_out/tab_aa.json > gpurun_out/tab_aa.txt; grep -E "ms/step|col_attention|our kernels" gpurun_out/tab_aa.txt; tail -2 gpurun_out/tab_aa.err
