#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_msg_reduce.py tests/test_gpu_ops.py tests/test_gpu_models.py -m gpu -q -x > gpurun_out/pytest_k.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_k.log | tail -6
for f in 0 1 0 1; do
  B200MP_UPD_BWD_SPLIT=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/split_$f.json 2>/dev/null; echo "== hi_medium, upd_bwd_split=$f"; python scripts/benchsum.py gpurun_out/split_$f.json | grep -E "ms/step|edge_update_bwd|message_bwd"
done
B200MP_UPD_BWD_SPLIT=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/split_small.json 2>/dev/null; python scripts/benchsum.py gpurun_out/split_small.json | grep -E "ms/step|edge_update_bwd"
