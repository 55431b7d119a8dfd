#!/bin/bash
# phase timeline of the SpMM tile kernel (experimental build exp_timeline.so): tools/gpu_timeline.sh <workload>...
for w in "$@"; do
  echo "== $w"
  SBLAS_B200_LIB=$PWD/sparsematrix_b200/lib/exp_timeline.so python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | tail -9
done
