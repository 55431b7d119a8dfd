#!/bin/bash
# One gpurun call: the default bench line (all five workloads) at N = $2 GPUs, plus the reference arm.
# Usage (repo root, under gpurun [--gpus N]):  bash tools/gpu_bench_all.sh <tag> [N] [extra bench args]
TAG=${1:-r02}; N=${2:-1}; shift; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
if [ "$N" = "1" ]; then
  python bench.py --gpus 1 "$@" > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench n1 exit $?"
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" \
      > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench n$N exit $?"
fi
tail -c 1500 $OUT/bench_n$N.err
python tools/benchfmt.py $OUT/bench_n$N.json
