#!/bin/bash
# Test infrastructure: compiles the reference's OWN negative sampler (its only native component,
# /root/reference/src/primitives/negative_sampling/negative_sampling.cpp) from where it lies into
# oracle/_ref/ (git-ignored; travels to the GPU box).  No reference source is copied into the repo.
# The reference builds it with cmake + pybind11; the same single translation unit compiles directly.
set -e
cd "$(dirname "$0")"
SRC=/root/reference/src/primitives/negative_sampling/negative_sampling.cpp
[ -f "$SRC" ] || { echo "reference source not present (GPU box): using the prebuilt oracle/_ref"; exit 0; }
mkdir -p _ref
EXT=$(python3 -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
g++ -O2 -shared -std=c++17 -fPIC $(python3 -m pybind11 --includes) "$SRC" -o _ref/negative_sampling$EXT
echo "built oracle/_ref/negative_sampling$EXT"
