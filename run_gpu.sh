mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
for v in default; do
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.log 2> gpurun_out/bench_$v.err; echo "bench $v rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$v.log').read().strip().splitlines()[-1])
print('$v', '%.3e'%d['value'], '%.3e'%d['e2e']['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rebuild_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'], 'step=%.1f'%d['ms_per_step'])
PY
done
B="python bench.py --steps 1 --warmup 1 --sites 2000000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"mh_update|csr_rebuild|suffstats_kernel" -s 24 -c 9 -o gpurun_out/prof_mh $B > gpurun_out/ncu2.log 2>&1
