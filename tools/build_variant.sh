#!/bin/bash
# build/lib_<name>.so from the working tree with extra nvcc flags: tools/build_variant.sh name [-DFOO=1 ...]
name=$1; shift
mkdir -p build
cd neuralstyletransferv1_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o ../../build/lib_$name.so nst_api.cu conv_tc.cu conv_direct.cu elementwise.cu postproc.cu masks.cu temporal.cu
