#!/bin/bash
# round 2, pass i: whole GPU suite, extra bench lines (tabgnn on hi_small, eth_lp, bf16), reference arm
mkdir -p gpurun_out; rm -f gpurun_out/parity_r02.txt
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^FAILED|passed|failed" gpurun_out/pytest_gpu.log | tail -12
grep -E "^bf16 |config 1 \(|edge-order|hub slice" gpurun_out/pytest_gpu.log | cut -c1-300 | head -12
timeout 900 python bench.py --model tabgnn --steps 5 --warmup 3 > gpurun_out/bench_tabgnn.json 2> gpurun_out/bench_tabgnn.err; echo "tabgnn rc=$?"; python scripts/benchsum.py gpurun_out/bench_tabgnn.json | head -40; tail -3 gpurun_out/bench_tabgnn.err
timeout 900 python bench.py --workload eth_lp --steps 20 --warmup 5 > gpurun_out/bench_eth_lp.json 2> gpurun_out/bench_eth_lp.err; echo "eth_lp rc=$?"; cut -c1-1500 gpurun_out/bench_eth_lp.json; tail -3 gpurun_out/bench_eth_lp.err
