#!/bin/bash
# last check of the round: whole GPU suite (-x, as the driver runs it), smoke, default bench line and reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; python scripts/benchsum.py gpurun_out/bench_default.json | head -1; wc -l gpurun_out/bench_default.json
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_default.json 2>/dev/null; echo "ref rc=$?"; head -c 200 gpurun_out/bench_ref_default.json
