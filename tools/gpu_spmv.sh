#!/bin/bash
# SpMV experiments on cfg4 (and cfg1): row order, lanes per row.
timeout 600 python -m pytest tests -m gpu -x -q -k "powerlaw or csr or config1 or spmv or golden" 2>&1 | tail -2
for spec in "0:2" "1:2" "1:1" "1:4" "1:8"; do
  so=${spec%%:*}; l=${spec#*:}
  SBLAS_B200_SORT_ROWS=$so SBLAS_B200_SPMV_LANES=$l timeout 300 python bench.py --workload cfg4 --steps 20 --warmup 3 --no-cpu-baseline > /tmp/b.json 2> /tmp/b.err
  python - <<PY
import json
try:
    d=json.load(open("/tmp/b.json")); r=d["roofline"]
    print("cfg4 sort $so lanes $l: step %.1f us kernel %.1f us hbm %.3f" % (d["ms_per_step"]*1e3, r["kernel_us"], r["frac"]))
except Exception as e:
    print("cfg4 $spec FAILED", e); print(open("/tmp/b.err").read()[-600:])
PY
done
python bench.py --workload cfg1 --steps 200 --warmup 10 --no-cpu-baseline | python -c "
import json,sys
d=json.load(sys.stdin); print('cfg1 step %.2f us warm %.2f us kernel %.2f us' % (d['ms_per_step']*1e3, d['warm_l2']['ms_per_step']*1e3, d['roofline']['kernel_us']))"
