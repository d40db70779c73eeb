#!/bin/bash
# Builds libaof_b200.so (sm_100a only) in-tree.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
SRC=$(ls csrc/*.cu)
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false \
   -Xcompiler -fPIC -Xptxas -v \
   -shared -o libaof_b200.so $SRC -lcudart_static -ldl -lrt -lpthread "$@"
