for v in base evl base evl; do
  LIB=""; [ "$v" != "base" ] && LIB=$PWD/sparsematrix_b200/lib/exp_$v.so
  SBLAS_B200_LIB=$LIB python bench.py --workload cfg4 --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.load(sys.stdin); print('cfg4 $v: step %.1f us kernel %.1f warm %.1f' % (d['ms_per_step']*1e3, d['roofline']['kernel_us'], d['warm_l2']['ms_per_step']*1e3))"
done
