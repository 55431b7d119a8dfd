#!/bin/bash
# Two GPUs: epilogue variants of the SpMM tile kernel with C in the peer's memory (SBLAS_B200_EPI).
OUT=gpurun_out/${1:-n2epi}
mkdir -p $OUT
timeout 500 python -m pytest tests/test_peer_output_gpu.py -x -q > $OUT/pytest_peer.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_peer.log
for w in cfg2 cfg3; do
 for epi in ${2:-auto 0 1 3}; do
  E=$epi; [ $epi = auto ] && E=""
  SBLAS_B200_EPI=$E timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
     bench.py --gpus 2 --workload $w --steps 30 --warmup 5 --no-cpu-baseline > $OUT/bench_${w}_epi$epi.json 2> $OUT/bench_${w}_epi$epi.err
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${w}_epi$epi.json"))
    print("$w epi=$epi N=2: local step %.1f us (warm %.1f) allgather %.1f us peer-store step %.1f us" % (d["ms_per_step"]*1e3, d["warm_l2"]["ms_per_step"]*1e3, d.get("reassembly_allgather_ms", -1)*1e3, d["reassembly_peer_store"]["ms_per_step"]*1e3))
except Exception as e: print("$w epi=$epi: no line", e); print(open("$OUT/bench_${w}_epi$epi.err").read()[-500:])
PY
 done
done
