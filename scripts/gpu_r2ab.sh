#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_col_attention.py -m gpu -q > gpurun_out/pytest_ab.log 2>&1; grep -E "^FAILED|passed|failed|Error" gpurun_out/pytest_ab.log | tail -8
for f in 0 1 0 1; do
  B200MP_ATT_LOOP=$f timeout 900 python bench.py --model tabgnn --steps 3 --warmup 2 --no-e2e --no-cpu-baseline > gpurun_out/tab_ab$f.json 2>/dev/null; echo "== att_loop=$f"; python scripts/benchsum.py gpurun_out/tab_ab$f.json | grep -E "ms/step|col_attention_fwd"
done
