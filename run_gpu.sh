mkdir -p gpurun_out
for w in bal8 testtree pairwise bal32; do
  python bench.py --workload $w --steps 1 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/ab_new_$w.log 2> gpurun_out/ab_new_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/ab_new_$w.log').read().strip().splitlines()[-1])
print('new $w', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'], 'rb_ms=%.3f'%d['roofline']['rebuild_kernel_avg_ms'])
PY
done
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"mh_update|csr_rebuild|suffstats_kernel" -s 13 -c 13 -o gpurun_out/prof_v4 $B > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
