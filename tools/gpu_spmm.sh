#!/bin/bash
# SpMM tile kernel experiments: parity tests, then cfg2 / cfg3 under variant overrides.
TAG=${1:-spmm}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_gpu.log
for v in "" e1 g1 72 82 g2 c1; do
  for w in cfg2 cfg3; do
    SBLAS_B200_TILE=$v timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline > $OUT/bench_${w}_v$v.json 2> $OUT/bench_${w}_v$v.err
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${w}_v$v.json")); r=d["roofline"]
    print("$w variant '$v': step %.1f us kernel %.1f us fp32 %.3f e2e %.1f" % (d["ms_per_step"]*1e3, r["kernel_us"], r["fp32_frac"], d["e2e"]["ms_per_step"]*1e3))
except Exception as e:
    print("$w variant '$v' FAILED", e); print(open("$OUT/bench_${w}_v$v.err").read()[-600:])
PY
  done
done
