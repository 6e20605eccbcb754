#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_msg_reduce.py tests/test_gpu_node_gemm.py tests/test_gpu_models.py -m gpu -q > gpurun_out/pytest_j.log 2>&1; grep -E "^FAILED|passed|failed" gpurun_out/pytest_j.log | tail -6
for f in 0 1; do
  B200MP_MSG_BWD64=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/mb64_$f.json 2>/dev/null; echo "== hi_medium, msg_bwd64=$f"; python scripts/benchsum.py gpurun_out/mb64_$f.json | grep -E "ms/step|message_bwd|edge_update_bwd"
  B200MP_MSG_BWD64=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --model gine > gpurun_out/mb64g_$f.json 2>/dev/null; echo "== gine hi_medium, msg_bwd64=$f"; python scripts/benchsum.py gpurun_out/mb64g_$f.json | grep -E "ms/step|message_bwd"
done
timeout 900 python -m pytest tests/test_gpu_parity_scale.py -m gpu -q -k "reverse" > gpurun_out/pytest_rev.log 2>&1; grep -E "^FAILED|passed|failed" gpurun_out/pytest_rev.log | tail -3
