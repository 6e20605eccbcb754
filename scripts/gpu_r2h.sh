#!/bin/bash
mkdir -p gpurun_out
run() { python scripts/benchsum.py $1 | grep -E "ms/step|message_fwd|edge_update|message_bwd|segment_reduce|node_linear64|node_pair|node_linear_wgrad"; }
for fl in 0 4096 1024; do
  B200MP_WGRAD_FLUSH=$fl timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/fl_$fl.json 2>/dev/null; echo "== hi_medium, edge-kernel drain interval $fl"; run gpurun_out/fl_$fl.json | grep -E "ms/step|_bwd"
done
(cd _r1 && timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > ../gpurun_out/r1_medium.json 2>/dev/null); echo "== round-1 tree, hi_medium"; run gpurun_out/r1_medium.json | grep -E "ms/step|_bwd|wgrad"
timeout 900 python -m pytest tests/test_bf16_storage.py tests/test_gpu_models.py tests/test_gpu_ops.py -m gpu -q -s > gpurun_out/pytest_bf16.log 2>&1; grep -E "^FAILED|passed|failed|^bf16" gpurun_out/pytest_bf16.log | tail -12
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --dtype bf16 > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "== bf16 storage, hi_medium"; python scripts/benchsum.py gpurun_out/bench_bf16.json | head -24; tail -3 gpurun_out/bench_bf16.err
