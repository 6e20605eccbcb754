#!/bin/bash
# partitioned bench under torchrun at N GPUs (bench only)
mkdir -p gpurun_out
N=${N:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
