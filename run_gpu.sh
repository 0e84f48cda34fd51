mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for v in base pf1 pf2; do
  if [ $v = base ]; then unset EPIEVO_B200_LIB; else export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so; fi
  for w in bal8 bal32 pairwise; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e --no-mcem > gpurun_out/ab_${v}_$w.log 2> gpurun_out/ab_${v}_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/ab_${v}_$w.log').read().strip().splitlines()[-1])
print('$v $w', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rb_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'])
PY
  done
done
unset EPIEVO_B200_LIB
python bench.py --steps 2 --warmup 2 > gpurun_out/bench_mcem.json 2> gpurun_out/bench_mcem.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_mcem.json').read().strip().splitlines()[-1])
print(d['mcem_iter'])
PY
