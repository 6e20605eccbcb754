#!/bin/bash
mkdir -p gpurun_out
for cfg in "1 0" "0 0" "1 1" "0 1"; do
  set -- $cfg
  rm -f gpurun_out/parity_r02.txt
  B200MP_TAB_STACK=$1 B200MP_EMB_LIB=$2 timeout 600 python -m pytest "tests/test_gpu_parity_scale.py::test_tab_models_h64_vs_oracle[tabgnn]" -m gpu -q > gpurun_out/pytest_w_$1$2.log 2>&1
  echo "== stack=$1 emb_lib=$2: $(tail -1 gpurun_out/pytest_w_$1$2.log)"
  grep -E "out0|grad.params|gnn_backbone.1.gnn_norm.module.weight|gnn_backbone.0.gnn_norm.module.weight" gpurun_out/parity_r02.txt | cut -c36-200
done
