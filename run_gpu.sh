mkdir -p gpurun_out
for w in bal8 testtree; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_$w.log 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; tail -3 gpurun_out/bench_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$w.log').read().strip().splitlines()[-1])
print('$w', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rebuild_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'], 'step=%.1f'%d['ms_per_step'], d['suffstats_pass'])
PY
done
