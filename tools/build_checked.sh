#!/bin/bash
# Builds build/librt_b200_checked.so: the library with -DRT_CHECKED, i.e. every index the kernels compute asserted against
# the size of its array (rt_device.h, RT_ASSERT).  Run the GPU suite through it with
#   RT_B200_LIB=$PWD/build/librt_b200_checked.so python -m pytest tests -m gpu -q
# (compute-sanitizer is closed on the GPU pool this was developed on; evidence: profiles/r02u_checked_build.txt).
set -e
cd "$(dirname "$0")/../raytrace_b200/csrc"
mkdir -p ../../build
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared -DRT_CHECKED \
  -o ../../build/librt_b200_checked.so rt_api.cpp rt_flatten.cpp rt_kernels.cu "$@"
