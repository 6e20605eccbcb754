#!/bin/bash
# 2-GPU pass at the end of round 2: partition parity tests (NCCL, real kernels), partitioned PNA and TABGNN bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_partition.py -m gpu -q -s > gpurun_out/pytest_partition.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED" gpurun_out/pytest_partition.log | tail -4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/n2_pna.json 2> gpurun_out/n2_pna.err; echo "== PNA N=2 rc=$?"
python scripts/benchsum.py gpurun_out/n2_pna.json 2>/dev/null | head -1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --model tabgnn --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/n2_tab.json 2> gpurun_out/n2_tab.err; echo "== TABGNN hi_small N=2 rc=$?"
python scripts/benchsum.py gpurun_out/n2_tab.json 2>/dev/null | head -1; tail -2 gpurun_out/n2_tab.err
