#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_ops.py tests/test_gpu_parity_scale.py tests/test_gpu_models.py tests/test_gpu_node_gemm.py -m gpu -q > gpurun_out/pytest_o.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_o.log | tail -12
