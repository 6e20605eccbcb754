#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_node_gemm.py tests/test_gpu_models.py tests/test_gpu_encoders.py -m gpu -q > gpurun_out/pytest_m.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_m.log | tail -12
timeout 300 python scripts/kbench_row_linear.py > gpurun_out/kbench_row_linear.txt 2>&1; cat gpurun_out/kbench_row_linear.txt | tail -12
timeout 600 python bench.py --model tabgnn --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/tab_m.json 2> gpurun_out/tab_m.err; python scripts/benchsum.py gpurun_out/tab_m.json | head -3; tail -2 gpurun_out/tab_m.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_tab.csv python bench.py --model tabgnn --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > /dev/null 2>&1; python scripts/summarize_launches.py gpurun_out/launches_tab.csv | head -40
