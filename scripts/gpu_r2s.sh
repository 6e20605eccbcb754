#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_zz_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/pytest_zz_alone.log
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_sampler.py tests/test_gpu_batch_inputs.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_zz_after.log 2>&1; echo "after rc=$?"; tail -3 gpurun_out/pytest_zz_after.log
