#!/bin/bash
# 2-GPU pass: NCCL partition parity (real kernels), partitioned bench lines (PNA hi_medium, TABGNN hi_small) with checksums
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_partition.py -m gpu -q -s > gpurun_out/pytest_partition.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_partition.log
grep -E "passed|failed|^FAILED|pytest rc" gpurun_out/pytest_partition.log | tail -5
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
python scripts/benchsum.py gpurun_out/bench_n2.json | head -3; grep -c "NCCL INFO" gpurun_out/bench_n2.err; grep -E "nranks|NVLS" gpurun_out/bench_n2.err | head -4
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n2.json') if l.startswith('{')][-1]); print("N=2 checksum", d['checksum']); print("e2e", d['e2e'].get('ms_per_step'))
PY
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --model tabgnn --workload hi_small --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_tabgnn_n2.json 2> gpurun_out/bench_tabgnn_n2.err; echo "tabgnn n2 rc=$?"
python scripts/benchsum.py gpurun_out/bench_tabgnn_n2.json | head -2; tail -3 gpurun_out/bench_tabgnn_n2.err
