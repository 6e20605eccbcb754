#!/bin/bash
# multi-GPU pass: NCCL partition parity test, then the partitioned bench under torchrun
mkdir -p gpurun_out
N=${N:-2}
timeout 600 python -m pytest tests/test_gpu_partition.py -m gpu -x -q > gpurun_out/pytest_partition.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_partition.log
tail -8 gpurun_out/pytest_partition.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -c 1800 gpurun_out/bench_n$N.json; tail -5 gpurun_out/bench_n$N.err
