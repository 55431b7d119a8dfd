#!/bin/sh
# build-b200.sh -- next to the reference's build-arm64.sh / build-armv7.sh / build-x86.sh:
# configure for NVIDIA B200 (sm_100a) and build the library, the static archive and the harness.
#   ./build-b200.sh [cuda-root]
set -e
cd "$(dirname "$0")"
CUDA_ROOT="${1:-${CUDA_HOME:-/usr/local/cuda}}"
make clean >/dev/null 2>&1 || true
./configure --target=B200 --cuda-root="$CUDA_ROOT"
make -j"$(nproc)" all
echo "built: lib/libsblas_b200.so lib/libsparsematrix.a bin/{sparse-matrix_test,parity_test,spmm_bench}"
