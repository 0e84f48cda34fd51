mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for w in bal8 testtree pairwise bal32; do
  python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || tail -3 gpurun_out/bench_$w.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
print('$w', '%.3e'%d['value'], 'e2e %.3e'%d['e2e']['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rb_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'], 'frac=%.4f'%d['roofline']['frac'], 'cpu %.3e'%d['cpu_baseline']['value'])
PY
done
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json | cut -c1-600
B="python bench.py --steps 1 --warmup 1 --burn-in 2 --batch 2 --no-cpu-baseline"
$B > gpurun_out/plain_l.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v5.csv $B > gpurun_out/ncu_l.log 2>&1
tail -1 gpurun_out/ncu_l.log | cut -c1-200
