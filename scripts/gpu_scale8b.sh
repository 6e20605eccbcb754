#!/bin/bash
# 8-GPU A/B: copy-engine all-gather off / on
mkdir -p gpurun_out
for ce in 0 1; do
  B200MP_CE_GATHER=$ce timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2953$ce bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n8_ce$ce.json 2> gpurun_out/n8_ce$ce.err; echo "== N=8 ce=$ce rc=$?"
  python scripts/benchsum.py gpurun_out/n8_ce$ce.json 2>/dev/null | grep -E "ms/step|edge_update_fwd|message_fwd|our kernels"; grep -E "b200mp\]" gpurun_out/n8_ce$ce.err | head -3
done
