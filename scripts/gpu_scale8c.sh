#!/bin/bash
# 8-GPU pass on the final tree of round 2: partitioned PNA hi_medium (the driver's SCALE line at N = 8) and config 5 (TABGNN, hi_medium x 8)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/n8_pna.json 2> gpurun_out/n8_pna.err; echo "== PNA N=8 rc=$?"
head -c 120 gpurun_out/n8_pna.json; echo; python scripts/benchsum.py gpurun_out/n8_pna.json 2>/dev/null | head -1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --model tabgnn --steps 2 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/n8_tab.json 2> gpurun_out/n8_tab.err; echo "== TABGNN hi_medium N=8 rc=$?"
python scripts/benchsum.py gpurun_out/n8_tab.json 2>/dev/null | head -1; tail -2 gpurun_out/n8_tab.err | cut -c1-200
