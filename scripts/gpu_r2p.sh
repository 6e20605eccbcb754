#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_ops.py tests/test_gpu_node_gemm.py tests/test_gpu_parity_scale.py tests/test_gpu_models.py -m gpu -q > gpurun_out/pytest_p.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_p.log | tail -12
B200MP_TC128=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small --hidden 128 > gpurun_out/h128_1.json 2> gpurun_out/h128_1.err; echo "== hi_small H=128, tc128=1"; python scripts/benchsum.py gpurun_out/h128_1.json > gpurun_out/h128_1.txt; cat gpurun_out/h128_1.txt; tail -2 gpurun_out/h128_1.err
