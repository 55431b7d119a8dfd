python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --workload cfg4 --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | tee gpurun_out/cfg4_base.json | python -c "
import json,sys; d=json.load(sys.stdin); print('cfg4: step %.1f us kernel %.1f warm %.1f frac %.3f' % (d['ms_per_step']*1e3, d['roofline']['kernel_us'], d['warm_l2']['ms_per_step']*1e3, d['roofline']['frac']))"
