#!/bin/bash
# Builds librt_b200.so (sm_100a only) in-tree; __graft_entry__.build() runs the same command.
set -e
cd "$(dirname "$0")/raytrace_b200/csrc"
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared \
  -o ../librt_b200.so rt_api.cpp rt_flatten.cpp rt_kernels.cu "$@"
