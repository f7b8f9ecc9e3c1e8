#!/bin/bash
# Builds the library of the last commit next to the working tree's, as kube_sqs_autoscaler_b200/libksareplay_head.so
# (KSA_LIBRARY=<that path> selects it): for A / B runs on one GPU box. The file is not tracked.
set -e
cd "$(dirname "$0")/.."
rm -rf build/head_csrc && mkdir -p build/head_csrc
for f in $(git ls-tree --name-only HEAD kube_sqs_autoscaler_b200/csrc/ | grep -E '\.(cu|cuh|h)$'); do git show HEAD:$f > build/head_csrc/$(basename $f); done
cd build/head_csrc
for f in *.cu; do nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --use_fast_math -c $f -o ${f%.cu}.o & done; wait
nvcc -shared -o ../../kube_sqs_autoscaler_b200/libksareplay_head.so *.o -cudart static -gencode arch=compute_100a,code=sm_100a
