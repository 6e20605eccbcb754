#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_msg_reduce.py tests/test_gpu_ops.py tests/test_bf16_storage.py -m gpu -q > gpurun_out/pytest_y.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_y.log | tail -8
for f in 0 1 0 1; do
  B200MP_SEG_TWOPASS=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/tp_$f.json 2>/dev/null; echo "== hi_medium, seg_twopass=$f"; python scripts/benchsum.py gpurun_out/tp_$f.json | grep -E "ms/step|segment_reduce"
done
B200MP_SEG_TWOPASS=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/tp_small.json 2>/dev/null; python scripts/benchsum.py gpurun_out/tp_small.json | grep -E "ms/step|segment_reduce"
