#!/bin/bash
# Multi-GPU round (gpurun --gpus N, NP=N; default 2): the peer-mapped output test, then bench lines at N GPUs.
TAG=${1:-n2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
[ -n "$SKIP_PYTEST" ] || timeout 500 python -m pytest tests/test_peer_output_gpu.py tests/test_sharding_gloo.py -x -q > $OUT/pytest_peer.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest_peer.log
for w in ${2:-cfg2 cfg3}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NP:-2} --master-addr 127.0.0.1 --master-port 29533 \
     bench.py --gpus ${NP:-2} --workload $w --steps 50 --warmup 5 --no-cpu-baseline > $OUT/bench_$w.json 2> $OUT/bench_$w.err
  echo "bench $w exit $?"; tail -3 $OUT/bench_$w.err; cut -c1-400 $OUT/bench_$w.json
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_$w.json"))
    print("$w N=${NP:-2}: value %.0f GFLOP/s step %.1f us allgather %.3f ms peer-store %s" % (d["value"], d["ms_per_step"]*1e3, d.get("reassembly_allgather_ms", -1), d.get("reassembly_peer_store")))
except Exception as e: print("no line", e)
PY
done
