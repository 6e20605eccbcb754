#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/n4_pna.json 2> gpurun_out/n4_pna.err; echo "== PNA N=4 rc=$?"
head -c 100 gpurun_out/n4_pna.json; echo; python scripts/benchsum.py gpurun_out/n4_pna.json 2>/dev/null | head -1
