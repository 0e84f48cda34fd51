mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
for v in base; do
  for w in bal8 bal32 testtree pairwise; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e --no-mcem > gpurun_out/ab_${v}_$w.log 2> gpurun_out/ab_${v}_$w.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/ab_${v}_$w.log').read().strip().splitlines()[-1])
    print('$v $w', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rb_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'])
except Exception as e:
    print('$v $w FAILED', open('gpurun_out/ab_${v}_$w.err').read()[-300:])
PY
  done
done
