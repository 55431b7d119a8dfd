#!/bin/sh
# build_ref.sh -- compile the UNMODIFIED reference hot path (two source files)
# from /root/reference into oracle/_ref/libsparsematrix_ref.so.
#
# Recipe = the reference's own generic build (configure:25 default CXXFLAGS,
# no --target, so the scalar AXPY body kernel.cc:568-582 is the one compiled)
# plus "-include cstdint" because kernel.cc uses uint8_t without including it
# (SURVEY.md appendix B).  The reference's own build system is not run; sources
# are read in place; outputs go only to oracle/_ref/ (git-ignored, but shipped
# to the GPU box with the snapshot).  -ffp-contract=off pins "no FMA", which is
# what the generic x86-64 build does anyway.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SBLAS_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -f "$REF/src/sparse/sparse-matrix.cc" ]; then
    echo "build_ref.sh: $REF not present; keeping prebuilt $OUT if any" >&2
    exit 0
fi
mkdir -p "$OUT"
CXXFLAGS="-std=c++11 -fPIC -O2 -fvisibility=hidden -ffp-contract=off -include cstdint -w"
g++ $CXXFLAGS -I"$REF/src/sparse" -c "$REF/src/sparse/kernel.cc" -o "$OUT/kernel.o"
g++ $CXXFLAGS -I"$REF/src/sparse" -c "$REF/src/sparse/sparse-matrix.cc" -o "$OUT/sparse-matrix.o"
g++ $CXXFLAGS -I"$REF/src/sparse" -c "$HERE/ref_shim.cc" -o "$OUT/ref_shim.o"
g++ -shared -o "$OUT/libsparsematrix_ref.so" "$OUT/kernel.o" "$OUT/sparse-matrix.o" "$OUT/ref_shim.o"
# the same sources instantiated with two-byte position deltas (ref_shim16.cc includes them), a library of its own
g++ $CXXFLAGS -I"$REF/src/sparse" -c "$HERE/ref_shim16.cc" -o "$OUT/ref_shim16.o"
g++ -shared -o "$OUT/libsparsematrix_ref16.so" "$OUT/ref_shim16.o"
rm -f "$OUT"/*.o
echo "built $OUT/libsparsematrix_ref.so $OUT/libsparsematrix_ref16.so"
