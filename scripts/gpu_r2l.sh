#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_node_gemm.py tests/test_gpu_msg_reduce.py -m gpu -q -x > gpurun_out/pytest_l.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_l.log | tail -6
for f in 0 1; do
  B200MP_TC128=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small --hidden 128 > gpurun_out/h128_$f.json 2> gpurun_out/h128_$f.err; echo "== hi_small H=128, tc128=$f"; python scripts/benchsum.py gpurun_out/h128_$f.json; tail -3 gpurun_out/h128_$f.err
done
B200MP_TC128=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small --hidden 128 --model gine > gpurun_out/h128g_1.json 2>/dev/null; echo "== gine hi_small H=128"; python scripts/benchsum.py gpurun_out/h128g_1.json
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/base_l.json 2>/dev/null; echo "== hi_medium H=64 (default path)"; python scripts/benchsum.py gpurun_out/base_l.json | head -12
timeout 900 python -m pytest tests/test_gpu_models.py tests/test_gpu_parity_scale.py -m gpu -q -x > gpurun_out/pytest_l2.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_l2.log | tail -6
