#!/bin/bash
# bench a workload with experimental library variants and tile overrides:
#   tools/gpu_variants.sh <workload> <lib variant[:TILE override]>...      (variant "base" = the product library)
W=$1; shift
for spec in "$@"; do
  v=${spec%%:*}; t=""; [[ "$spec" == *:* ]] && t=${spec#*:}
  LIB=""; [ "$v" != "base" ] && LIB=$PWD/sparsematrix_b200/lib/exp_$v.so
  SBLAS_B200_TILE=$t SBLAS_B200_LIB=$LIB timeout 300 python bench.py --workload $W --steps 30 --warmup 5 --no-cpu-baseline > /tmp/b.json 2> /tmp/b.err
  python - <<PY
import json
try:
    d=json.load(open("/tmp/b.json")); r=d["roofline"]
    print("$W variant $spec: step %.1f us kernel %.1f us fp32 %.3f" % (d["ms_per_step"]*1e3, r["kernel_us"], r["fp32_frac"]))
except Exception as e:
    print("$W variant $spec FAILED", e); print(open("/tmp/b.err").read()[-600:])
PY
done
