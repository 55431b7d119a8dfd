python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --no-cpu-baseline > gpurun_out/final_cfg2.json 2>gpurun_out/final_cfg2.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/final_cfg2.json')); print('cfg2: step %.1f us e2e %.0f value %.0f launches %s clocks %s' % (d['ms_per_step']*1e3, d['e2e']['value'], d['value'], d['gpu_launches'], d['clocks']))"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
