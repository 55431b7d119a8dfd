#!/bin/sh
# build_ref_callers.sh -- the reference's OWN callers, compiled UNMODIFIED from /root/reference where they lie,
# against THIS build's headers (sparsematrix_b200/src/sparse/{sparse-matrix.h,kernel.h}: the reference's names)
# and archive (sparsematrix_b200/lib/libsparsematrix.a, the reference's library name, src/sparse/Makefile:9):
#   src/sparse/sparse-matrix_test.cc:4-13   -> _ref/callers/sparse-matrix_test   (SelfTest through the GPU path)
#   src/sparse/kernel_test.cc:27-57         -> _ref/callers/kernel_test          (sblas_trans_kernel on the device-backed functions)
#   src/test/blas_test.cc:31-47 + blas_test.h -> _ref/callers/blas_test           (the harness: sparse row computed on the B200,
#                                              checked by its own 10 % rule against cblas_sgemm of a stock OpenBLAS)
# This is the drop-in proof of SURVEY.md 8(b) "unchanged call sites must work unmodified".  Nothing of the reference is
# copied into the repo; outputs go to oracle/_ref/ (git-ignored, shipped to the GPU box).  tests/test_reference_callers_gpu.py
# runs the binaries.  OpenBLAS: the x86-64 library bundled with a Python wheel of this image (unprefixed cblas_*, LP64);
# headers: the reference's own openblas_arm64_plus/include/cblas.h.  No OpenBLAS found -> blas_test is skipped.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
REF="${SBLAS_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref/callers"
PKG="$ROOT/sparsematrix_b200"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
if [ ! -f "$REF/src/sparse/sparse-matrix_test.cc" ]; then
    echo "build_ref_callers.sh: $REF not present; keeping prebuilt $OUT if any" >&2
    exit 0
fi
if [ ! -f "$PKG/lib/libsparsematrix.a" ]; then
    echo "build_ref_callers.sh: build sparsematrix_b200 first" >&2
    exit 1
fi
mkdir -p "$OUT"
CXXFLAGS="-std=c++11 -O2 -include cstdint -w"
LINK="$NVCC -gencode arch=compute_100a,code=sm_100a -cudart static"
g++ $CXXFLAGS -I"$PKG/src/sparse" -c "$REF/src/sparse/sparse-matrix_test.cc" -o "$OUT/sparse-matrix_test.o"
$LINK -o "$OUT/sparse-matrix_test" "$OUT/sparse-matrix_test.o" "$PKG/lib/libsparsematrix.a"
g++ $CXXFLAGS -I"$PKG/src/sparse" -c "$REF/src/sparse/kernel_test.cc" -o "$OUT/kernel_test.o"
$LINK -o "$OUT/kernel_test" "$OUT/kernel_test.o" "$PKG/lib/libsparsematrix.a"
BLAS="$(python3 - <<'PY'
import glob, sysconfig
hits = sorted(glob.glob(sysconfig.get_paths()["purelib"] + "/opencv_python_headless.libs/libopenblasp-*.so"))
print(hits[0] if hits else "")
PY
)"
if [ -n "$BLAS" ]; then
    # -I order: this build's sparse-matrix.h first; blas_test.h's "../utils/dev-tools.h" resolves inside the reference tree
    g++ $CXXFLAGS -include stdint.h -I"$PKG/src/sparse" -I"$REF/openblas_arm64_plus/include" -I"$REF/src/test" \
        -c "$REF/src/test/blas_test.cc" -o "$OUT/blas_test.o"
    gcc -O2 -c "$HERE/openblas_stub.c" -o "$OUT/openblas_stub.o"
    $LINK -o "$OUT/blas_test" "$OUT/blas_test.o" "$OUT/openblas_stub.o" "$PKG/lib/libsparsematrix.a" "$BLAS" \
        -Xlinker -rpath -Xlinker "$(dirname "$BLAS")" -Xlinker --allow-shlib-undefined
    echo "$(dirname "$BLAS")" > "$OUT/blas_libdir.txt"
else
    echo "build_ref_callers.sh: no x86-64 OpenBLAS in this image; blas_test not built" >&2
    rm -f "$OUT/blas_test"
fi
rm -f "$OUT"/*.o
echo "built $(ls "$OUT" | tr '\n' ' ')"
