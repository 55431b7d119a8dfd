#!/bin/bash
# One gpurun call: GPU parity tests, the C++ harness, bench lines for every config, the reference arm,
# and an ncu launch list of the default bench (full captures: tools/gpu_ncu.sh, one per gpurun call).
# Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh [tag]
TAG=${1:-r02}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/gpu.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?" | tee -a $OUT/smoke.log
(cd sparsematrix_b200 && ./bin/sparse-matrix_test && ./bin/parity_test) > $OUT/cxx_harness.log 2>&1; echo "cxx exit $?" | tee -a $OUT/cxx_harness.log
for w in cfg2 cfg1 cfg3 cfg4 cfg5; do
  python bench.py --workload $w > $OUT/bench_$w.json 2> $OUT/bench_$w.err; echo "bench $w exit $?"
done
python bench.py --impl reference --steps 5 --warmup 3 > $OUT/bench_ref_cfg2.json 2> $OUT/bench_ref_cfg2.err
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_cfg2.csv $CMD > $OUT/ncu_launches.log 2>&1
tail -3 $OUT/pytest_gpu.log; cat $OUT/bench_cfg2.json
