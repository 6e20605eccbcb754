#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_msg_reduce.py tests/test_gpu_ops.py tests/test_gpu_models.py -m gpu -q > gpurun_out/pytest_z.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_z.log | tail -8
for f in 0 1 0 1; do
  B200MP_L2PF=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/l2pf_$f.json 2>/dev/null; echo "== hi_medium, l2pf=$f"; python scripts/benchsum.py gpurun_out/l2pf_$f.json | grep -E "ms/step|message_fwd|message_bwd"
done
for f in 0 1; do
  B200MP_L2PF=$f timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --model gine > gpurun_out/l2pfg_$f.json 2>/dev/null; echo "== gine hi_medium, l2pf=$f"; python scripts/benchsum.py gpurun_out/l2pfg_$f.json | grep -E "ms/step|message_fwd|message_bwd"
done
