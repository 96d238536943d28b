#!/bin/bash
# build one libb200pt.so variant with extra compile-time switches:  bash benchmarks/build_variant.sh NAME -DPT_X=1 ...
name=$1; shift
mkdir -p variants/$name
/usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Iinclude "$@" \
  -shared -cudart static -o variants/$name/libb200pt.so cuda_pathtracer_b200/csrc/pt_kernels.cu cuda_pathtracer_b200/csrc/pt_api.cu \
  cuda_pathtracer_b200/csrc/pt_gbvh.cu -Lcuda_pathtracer_b200 -lb200sc -Xlinker -rpath -Xlinker '$ORIGIN' 2>&1 | grep -i "error" ; echo "built $name"
