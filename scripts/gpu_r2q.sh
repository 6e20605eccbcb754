#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_msg_reduce.py tests/test_gpu_ops.py -m gpu -q > gpurun_out/pytest_q.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_q.log | tail -8
for f in 0 1 0 1; do
  B200MP_UPD_GPF=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/gpf_$f.json 2>/dev/null; echo "== hi_medium, upd_gpf=$f"; python scripts/benchsum.py gpurun_out/gpf_$f.json | grep -E "ms/step|edge_update_fwd"
done
B200MP_UPD_GPF=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/gpf_small.json 2>/dev/null; python scripts/benchsum.py gpurun_out/gpf_small.json | grep -E "ms/step|edge_update_fwd"
