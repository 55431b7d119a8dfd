#!/bin/bash
# e2e (host buffers) step time: overlapped host path off / on / sliced with different slice heights.
# Usage (under gpurun):  bash tools/gpu_e2e.sh [tag]
TAG=${1:-e2e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; tail -3 $OUT/pytest_gpu.log
run() {  # workload label env...
  w=$1; label=$2; shift 2
  env "$@" python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > $OUT/bench_${w}_$label.json 2> $OUT/bench_${w}_$label.err
  python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_${w}_$label.json"))
    print("$w $label: step %.1f us  e2e %.1f us (%.1f GFLOP/s)" % (d["ms_per_step"]*1e3, d["e2e"]["ms_per_step"]*1e3, d["e2e"]["value"]))
except Exception as e:
    print("$w $label: failed", e)
PY
}
run cfg2 off SBLAS_B200_HOST_OVERLAP=0
run cfg2 auto X=1
run cfg2 s0 SBLAS_B200_HOST_SLICE_ROWS=0
run cfg2 s64 SBLAS_B200_HOST_SLICE_ROWS=64
run cfg3 off SBLAS_B200_HOST_OVERLAP=0
run cfg3 auto X=1
run cfg3 s0 SBLAS_B200_HOST_SLICE_ROWS=0
run cfg3 s64 SBLAS_B200_HOST_SLICE_ROWS=64
run cfg3 s128 SBLAS_B200_HOST_SLICE_ROWS=128
run cfg3 s256 SBLAS_B200_HOST_SLICE_ROWS=256
run cfg1 auto X=1
run cfg4 auto X=1
