mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for w in bal8 bal32; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_$w.log 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$w.log').read().strip().splitlines()[-1])
print('$w', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rebuild_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'], 'step=%.1f'%d['ms_per_step'], 'acc=%.4f'%d['acc_rate'])
PY
done
B="python bench.py --steps 1 --warmup 1 --sites 2000000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"suffstats_kernel" -s 1 -c 1 -o gpurun_out/prof_ss $B > gpurun_out/ncu3.log 2>&1
