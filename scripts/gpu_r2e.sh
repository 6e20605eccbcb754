#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/debug_r2.py > gpurun_out/debug_r2e.log 2>&1; grep -E "^wgrad|^linear_group" gpurun_out/debug_r2e.log
for cfg in "1 8192" "0 8192" "1 0" "0 0"; do set -- $cfg
  B200MP_CARVEOUT=$1 B200MP_WGRAD_FLUSH=$2 timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/cv_$1_$2.json 2>/dev/null
  echo "carveout=$1 flush=$2"; python scripts/benchsum.py gpurun_out/cv_$1_$2.json | grep -E "ms/step|message_fwd|edge_update|message_bwd|node_linear64"
done
B200MP_UPD_PIPE=0 timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/pipe0.json 2>/dev/null; echo "pipe=0"; python scripts/benchsum.py gpurun_out/pipe0.json | grep -E "ms/step|edge_update_fwd"
timeout 1500 python -m pytest tests/test_gpu_parity_scale.py tests/test_readout.py tests/test_gpu_msg_reduce.py tests/test_gpu_models.py -m gpu -q -s > gpurun_out/pytest_sel.log 2>&1; grep -E "^FAILED|passed|failed" gpurun_out/pytest_sel.log | tail
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_e.json 2>/dev/null; python scripts/benchsum.py gpurun_out/bench_e.json | head -30
