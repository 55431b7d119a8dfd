#!/bin/bash
# Build an experimental variant of the library: tools/build_variant.sh NAME "<extra nvcc flags>" [sources...]
# -> sparsematrix_b200/lib/exp_NAME.so (select with SBLAS_B200_LIB=...).  Experiments only.  The named sources (default:
# spmm_tile.cu and product.cu) are recompiled with the extra flags, every other object is the regular build's.
set -e
cd "$(dirname "$0")/../sparsematrix_b200"
NAME=$1; FLAGS=$2; shift; shift
SRCS=${@:-src/sparse/spmm_tile.cu src/sparse/product.cu}
NV="nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -I../include -Isrc/utils -Isrc/sparse"
mkdir -p /tmp/exp_$NAME; rm -f /tmp/exp_$NAME/*.o
SKIP=""
for f in $SRCS; do $NV $FLAGS -Xptxas -v -c $f -o /tmp/exp_$NAME/$(basename $f .cu).o 2>&1 | grep -E "error|spill" | grep -v " 0 bytes spill" || true; SKIP="$SKIP ${f%.cu}.o"; done
OTHERS=""
for o in src/utils/*.o src/sparse/*.o; do case " $SKIP " in *" $o "*) ;; *) OTHERS="$OTHERS $o";; esac; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o lib/exp_$NAME.so /tmp/exp_$NAME/*.o $OTHERS
echo built lib/exp_$NAME.so
