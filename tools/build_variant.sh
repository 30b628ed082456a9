#!/bin/bash
# Builds another copy of the library with extra nvcc flags, for A/B runs through GT_LIB inside one GPU session:
#   tools/build_variant.sh ab/base.so -DGT_SMALL_SMEM=0
out=$1; shift
mkdir -p "$(dirname "$out")"
d=gpu-topology-on-k8s_b200/csrc
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared "$@" \
     -o "$out" $d/api.cu $d/host.cc $d/extender.cc $d/gaia_tree.cc $d/wire.cc -ldl
