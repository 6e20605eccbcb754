#!/bin/bash
# 8-GPU pass: partitioned bench (hi_medium) under a few NCCL settings, then the TABGNN config-5 line
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n8_$tag.json 2> gpurun_out/n8_$tag.err; echo "== $tag rc=$?"; python scripts/benchsum.py gpurun_out/n8_$tag.json 2>/dev/null | head -1; }
run default A=1
run nchan32 NCCL_MIN_NCHANNELS=32
run ring NCCL_ALGO=Ring
run nvls NCCL_ALGO=NVLS
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --model tabgnn --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n8_tabgnn.json 2> gpurun_out/n8_tabgnn.err; echo "== tabgnn hi_medium x8 rc=$?"; python scripts/benchsum.py gpurun_out/n8_tabgnn.json 2>/dev/null | head -1; tail -2 gpurun_out/n8_tabgnn.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/n8_full.json 2> gpurun_out/n8_full.err; echo "== full line rc=$?"; python scripts/benchsum.py gpurun_out/n8_full.json 2>/dev/null | head -1
