#!/bin/bash
# config 5 on the final tree: TABGNN, hi_medium over 8 x B200
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --model tabgnn --steps 2 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/n8_tab.json 2> gpurun_out/n8_tab.err; echo "== TABGNN hi_medium N=8 rc=$?"
python scripts/benchsum.py gpurun_out/n8_tab.json 2>/dev/null | head -1
