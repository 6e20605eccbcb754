#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_column_stack.py tests/test_gpu_col_attention.py tests/test_gpu_row_norm.py tests/test_gpu_models.py tests/test_gpu_parity_scale.py -m gpu -q > gpurun_out/pytest_v.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_v.log | tail -8
B200MP_TORCH_PROFILE=gpurun_out/tab_torch_profile.txt timeout 900 python bench.py --model tabgnn --steps 2 --warmup 2 --no-e2e --no-cpu-baseline > gpurun_out/tab_v.json 2> gpurun_out/tab_v.err; python scripts/benchsum.py gpurun_out/tab_v.json > gpurun_out/tab_v.txt; cat gpurun_out/tab_v.txt; tail -3 gpurun_out/tab_v.err
grep -E "aten::|Memcpy|sgemm" gpurun_out/tab_torch_profile.txt | awk '{n=$1; print substr(n,1,60), $(NF-4), $(NF-3), $(NF)}' | head -14
