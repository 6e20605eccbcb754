#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_node_gemm.py tests/test_gpu_col_attention.py tests/test_gpu_models.py tests/test_gpu_encoders.py -m gpu -q > gpurun_out/pytest_u.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_u.log | tail -8
B200MP_TORCH_PROFILE=gpurun_out/tab_torch_profile.txt timeout 900 python bench.py --model tabgnn --steps 2 --warmup 2 --no-e2e --no-cpu-baseline > gpurun_out/tab_u.json 2> gpurun_out/tab_u.err; python scripts/benchsum.py gpurun_out/tab_u.json > gpurun_out/tab_u.txt; head -3 gpurun_out/tab_u.txt; tail -3 gpurun_out/tab_u.err
grep -E "aten::|Memcpy|sgemm" gpurun_out/tab_torch_profile.txt | cut -c1-60,150-200 | head -20
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/base_u.json 2>/dev/null; python scripts/benchsum.py gpurun_out/base_u.json | head -1
