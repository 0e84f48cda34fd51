mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for v in default mb5 mb6 mb8; do
  if [ $v = default ]; then unset EPIEVO_B200_LIB; else export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so; fi
  python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_$v.log 2> gpurun_out/bench_$v.err; echo "bench $v rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$v.log').read().strip().splitlines()[-1])
print('$v', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rebuild_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'], 'step=%.1f'%d['ms_per_step'])
PY
done
