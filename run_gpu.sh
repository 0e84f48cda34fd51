mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
for v in default bin48 nopipe b3; do
  if [ $v = default ]; then unset EPIEVO_B200_LIB; else export EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_$v.so; fi
  for w in bal8 bal32 testtree; do
    timeout 300 $B --workload $w > gpurun_out/ab_${v}_$w.json 2> gpurun_out/ab_${v}_$w.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ab_${v}_$w.json").read().strip().splitlines()[-1])
    print("$v $w", d["suffstats_pass"]["ms"], d["roofline"]["avg_launch_ms"], d["roofline"]["rebuild_kernel_avg_ms"])
except Exception as e: print("$v $w FAILED", e)
PY
  done
done
unset EPIEVO_B200_LIB
timeout 600 ncu --set full --clock-control none --import-source on -k regex:suffstats_tile -s 2 -c 2 -o gpurun_out/prof_ss_tile7 $B > gpurun_out/ncu_ss.log 2>&1
tail -2 gpurun_out/ncu_ss.log
