#!/bin/bash
# SpMM tile kernel experiments: parity tests, then workloads under variant overrides.
# usage: tools/gpu_spmm.sh <tag> "<variants for cfg2>" "<variants for cfg3>"
TAG=${1:-spmm}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_gpu.log
run() {
  w=$1; v=$2
  SBLAS_B200_TILE=$v timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > $OUT/bench_${w}_v$v.json 2> $OUT/bench_${w}_v$v.err
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${w}_v$v.json")); r=d["roofline"]
    print("$w variant '$v': step %.1f us (warm %.1f) kernel %.1f us fp32 %.3f e2e %.1f" % (d["ms_per_step"]*1e3, d["warm_l2"]["ms_per_step"]*1e3, r["kernel_us"], r["fp32_frac"], d["e2e"]["ms_per_step"]*1e3))
except Exception as e:
    print("$w variant '$v' FAILED", e); print(open("$OUT/bench_${w}_v$v.err").read()[-600:])
PY
}
for v in $2; do run cfg2 $v; done
for v in $3; do run cfg3 $v; done
