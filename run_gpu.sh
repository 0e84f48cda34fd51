mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -4 gpurun_out/pytest.log
B="python bench.py --steps 2 --warmup 1 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline --no-mcem"
for v in default noq default noq; do
  if [ $v = default ]; then unset EPIEVO_B200_LIB; else export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so; fi
  for w in bal8 testtree; do
    timeout 300 $B --workload $w > gpurun_out/ab_${v}_$w.json 2> gpurun_out/ab_${v}_$w.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ab_${v}_$w.json").read().strip().splitlines()[-1])
    print("$v $w", d["suffstats_pass"]["ms"], d["roofline"]["avg_launch_ms"], d["roofline"]["rebuild_kernel_avg_ms"], d["acc_rate"])
except Exception as e: print("$v $w FAILED", e)
PY
  done
done
unset EPIEVO_B200_LIB
timeout 500 python bench.py --workload testtree --sites 100000000 --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem > gpurun_out/jd_c3_1e8.json 2> gpurun_out/jd_c3_1e8.err; tail -c 400 gpurun_out/jd_c3_1e8.json; tail -3 gpurun_out/jd_c3_1e8.err
