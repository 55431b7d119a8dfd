#!/bin/bash
# bench a workload with each experimental library variant: tools/gpu_variants.sh <workload> <variant>...
W=$1; shift
for v in "$@"; do
  LIB=""; [ "$v" != "base" ] && LIB=$PWD/sparsematrix_b200/lib/exp_$v.so
  SBLAS_B200_LIB=$LIB timeout 300 python bench.py --workload $W --steps 30 --warmup 5 --no-cpu-baseline > /tmp/b.json 2> /tmp/b.err
  python - <<PY
import json
try:
    d=json.load(open("/tmp/b.json")); r=d["roofline"]
    print("$W variant $v: step %.1f us kernel %.1f us fp32 %.3f" % (d["ms_per_step"]*1e3, r["kernel_us"], r["fp32_frac"]))
except Exception as e:
    print("$W variant $v FAILED", e); print(open("/tmp/b.err").read()[-600:])
PY
done
