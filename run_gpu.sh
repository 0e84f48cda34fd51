mkdir -p gpurun_out
for v in mb4 mb3; do
  export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so
  python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_$v.log 2> gpurun_out/bench_$v.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$v.log').read().strip().splitlines()[-1])
print('$v', '%.3e'%d['value'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'])
PY
done
unset EPIEVO_B200_LIB
B="python bench.py --steps 1 --warmup 1 --sites 2000000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"mh_update" -s 12 -c 2 -o gpurun_out/prof_mh $B > gpurun_out/ncu2.log 2>&1
