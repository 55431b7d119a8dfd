#!/bin/bash
# compute-sanitizer over the C++ parity harness (every product kernel family, the device CopyForm
# route, the pageable host path).  memcheck over the built-in sweep minus the two large cases,
# then racecheck + synccheck on one SpMM tile case and one SpMV case.
# Usage (under gpurun, from the repo root):  bash tools/gpu_sanitize.sh [tag]
TAG=${1:-sanitize}
OUT=gpurun_out/$TAG
mkdir -p $OUT
BIN=sparsematrix_b200/bin/parity_test
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # name, tool, args...
  name=$1; tool=$2; shift 2
  timeout 600 $CS --tool $tool --error-exitcode 9 --print-limit 20 $BIN "$@" > $OUT/${name}.log 2>&1
  echo "$name ($tool $*): exit $? -- $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $OUT/${name}.log | tail -1)"
}
run mem_spmv      memcheck  1   1023 2047 0.25
run mem_gather    memcheck  9   700  333  0.02
run mem_tile16    memcheck  7   1023 2047 0.25
run mem_tile      memcheck  117 1023 2047 0.25
run mem_tile_big  memcheck  128 2048 1024 0.10
run race_tile     racecheck 117 1023 2047 0.25
run race_spmv     racecheck 1   1023 2047 0.25
run sync_tile     synccheck 117 1023 2047 0.25
run init_tile     initcheck 37  700  333  0.10
