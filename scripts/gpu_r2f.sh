#!/bin/bash
# same-box A/B against the round-1 tree (_r1/), hi_small and hi_medium, plus pipe on/off at hi_medium and the selected tests
mkdir -p gpurun_out
run() { python scripts/benchsum.py $1 | grep -E "ms/step|message_fwd|edge_update|message_bwd|segment_reduce|node_linear64|node_pair|node_linear_wgrad"; }
(cd _r1 && timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > ../gpurun_out/r1_small.json 2>/dev/null); echo "== round-1 tree, hi_small"; run gpurun_out/r1_small.json
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --workload hi_small > gpurun_out/r2_small.json 2>/dev/null; echo "== this tree, hi_small"; run gpurun_out/r2_small.json
(cd _r1 && timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > ../gpurun_out/r1_medium.json 2>/dev/null); echo "== round-1 tree, hi_medium"; run gpurun_out/r1_medium.json
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2_medium.json 2>/dev/null; echo "== this tree, hi_medium"; run gpurun_out/r2_medium.json
B200MP_UPD_PIPE=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2_medium_pipe0.json 2>/dev/null; echo "== this tree, hi_medium, pipe=0"; run gpurun_out/r2_medium_pipe0.json | grep -E "ms/step|edge_update_fwd"
timeout 1500 python -m pytest tests/test_gpu_parity_scale.py tests/test_gpu_node_gemm.py -m gpu -q -s > gpurun_out/pytest_sel.log 2>&1; grep -E "^FAILED|passed|failed" gpurun_out/pytest_sel.log | tail -5
