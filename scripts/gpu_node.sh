#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_node_gemm.py -m gpu -x -q -s > gpurun_out/node_test.log 2>&1; echo "rc=$?" >> gpurun_out/node_test.log
tail -25 gpurun_out/node_test.log
KB_ARGS="--only node" ; timeout 300 python scripts/kbench.py --only "node" > gpurun_out/kbench_node.log 2>&1; timeout 300 python scripts/kbench.py --only "mm3" >> gpurun_out/kbench_node.log 2>&1; grep " ms" gpurun_out/kbench_node.log
