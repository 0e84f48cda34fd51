mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sampler.py tests/test_gpu_golden.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
for w in bal8 bal32 testtree; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e --no-mcem > gpurun_out/ab_ss_$w.log 2> gpurun_out/ab_ss_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/ab_ss_$w.log').read().strip().splitlines()[-1])
print('$w', '%.3e'%d['value'], 'ss_ms=%.3f'%d['suffstats_pass']['ms'], 'frac=%.3f'%d['suffstats_pass']['frac_of_hbm_peak'])
PY
done
