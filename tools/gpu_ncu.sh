#!/bin/bash
# One ncu --set full capture of a kernel of a bench workload.  Usage: tools/gpu_ncu.sh <tag> <workload> <kernel regex> [skip]
TAG=$1; W=$2; K=$3; SKIP=${4:-3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python bench.py --workload $W --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/plain_$W.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 2 -o $OUT/prof_$W $CMD > $OUT/ncu_$W.log 2>&1
tail -3 $OUT/ncu_$W.log | cut -c1-300
