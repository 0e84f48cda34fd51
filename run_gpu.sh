mkdir -p gpurun_out
for v in g1 g2 g3; do
  export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so
  for w in bal8 bal32 testtree; do
  python bench.py --workload $w --steps 1 --warmup 1 --burn-in 2 --batch 2 --no-cpu-baseline --no-e2e --no-mcem > gpurun_out/ab_${v}_$w.log 2> gpurun_out/ab_${v}_$w.err; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/ab_${v}_$w.log').read().strip().splitlines()[-1])
    print('$v $w', 'ss_ms=%.3f'%d['suffstats_pass']['ms'])
except Exception as e:
    print('$v $w FAILED', open('gpurun_out/ab_${v}_$w.err').read()[-300:])
PY
  done
done
