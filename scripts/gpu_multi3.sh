#!/bin/bash
# 2-GPU pass: copy-engine all-gather on/off -- parity test and partitioned bench
mkdir -p gpurun_out
B200MP_CE_GATHER=1 timeout 900 python -m pytest tests/test_gpu_partition.py -m gpu -q -s -k "pna or gine" > gpurun_out/pytest_partition_ce.log 2>&1; echo "pytest(ce) rc=$?"
grep -E "passed|failed|^FAILED|b200mp\]" gpurun_out/pytest_partition_ce.log | tail -6
for ce in 0 1; do
  B200MP_CE_GATHER=$ce timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$ce bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n2_ce$ce.json 2> gpurun_out/n2_ce$ce.err; echo "== N=2 ce=$ce rc=$?"
  python scripts/benchsum.py gpurun_out/n2_ce$ce.json 2>/dev/null | grep -E "ms/step|edge_update_fwd|message_fwd" ; grep -E "b200mp\]|Error|error" gpurun_out/n2_ce$ce.err | head -5
done
