#!/bin/bash
# Build an experimental variant of the library: tools/build_variant.sh NAME "<extra nvcc flags>"
# -> sparsematrix_b200/lib/exp_NAME.so (select with SBLAS_B200_LIB=...).  Experiments only.
set -e
cd "$(dirname "$0")/../sparsematrix_b200"
NAME=$1; FLAGS=$2
NV="nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -I../include -Isrc/utils -Isrc/sparse"
mkdir -p /tmp/exp_$NAME
for f in src/sparse/spmm_tile.cu src/sparse/product.cu; do $NV $FLAGS -Xptxas -v -c $f -o /tmp/exp_$NAME/$(basename $f .cu).o; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o lib/exp_$NAME.so /tmp/exp_$NAME/*.o src/utils/device_format.o src/utils/tile_format.o src/utils/copyform.o src/utils/peer_memory.o src/utils/pinned_memory.o src/utils/device_registry.o src/sparse/sparse-matrix.o src/sparse/host_capi.o src/sparse/spmv_dense.o src/sparse/kernel_shim.o src/sparse/kernel.o
echo built lib/exp_$NAME.so
