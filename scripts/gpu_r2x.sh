#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_column_stack.py tests/test_gpu_node_gemm.py tests/test_gpu_models.py tests/test_gpu_parity_scale.py tests/test_gpu_partition.py -m gpu -q > gpurun_out/pytest_x.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_x.log | tail -8
timeout 900 python bench.py --model tabgnn --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/tab_x.json 2> gpurun_out/tab_x.err; python scripts/benchsum.py gpurun_out/tab_x.json > gpurun_out/tab_x.txt; cat gpurun_out/tab_x.txt; tail -3 gpurun_out/tab_x.err
