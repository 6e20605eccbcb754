#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_umma.py -m gpu -q -s > gpurun_out/umma_test.log 2>&1; echo "rc=$?" >> gpurun_out/umma_test.log
tail -30 gpurun_out/umma_test.log
