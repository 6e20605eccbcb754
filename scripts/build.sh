#!/bin/bash
# in-tree build + register / spill summary of the tcgen05 kernels
cd "$(dirname "$0")/.."
python models-for-relational-multimodal-data_b200/build.py > /tmp/build.log 2>&1 || { tail -40 /tmp/build.log; exit 1; }
grep -h "registers\|spill\|Compiling entry.*umma" models-for-relational-multimodal-data_b200/build/edge_*umma.o.log | grep -A2 "umma_kernel" | grep -v "^--" | cut -c1-150
