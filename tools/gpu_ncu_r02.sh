#!/bin/bash
# r02 evidence: ncu --set full captures of the dominant kernels and per-launch time lists (one gpurun call, one GPU).
OUT=gpurun_out/${NCU_TAG:-r02n}
mkdir -p $OUT
for spec in "cfg3:spmm_tc_kernel:3" "cfg4:spmv_blocked:6" "cfg2:spmm_tc_kernel:3"; do
  W=${spec%%:*}; rest=${spec#*:}; K=${rest%%:*}; SKIP=${rest#*:}
  CMD="python bench.py --workload $W --steps 3 --warmup 3 --no-cpu-baseline"
  $CMD > $OUT/plain_$W.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 2 -o $OUT/prof_$W $CMD > $OUT/ncu_$W.log 2>&1
  tail -2 $OUT/ncu_$W.log | cut -c1-200
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$W.csv $CMD > $OUT/ncu_launches_$W.log 2>&1
done
