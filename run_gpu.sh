mkdir -p gpurun_out
for v in noskip oldcall_skip; do
  export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so
  for w in bal8 testtree; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/ab_${v}_$w.log 2> gpurun_out/ab_${v}_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/ab_${v}_$w.log').read().strip().splitlines()[-1])
print('$v $w', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rb_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'])
PY
  done
done
