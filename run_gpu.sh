mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_gpu.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['roofline']['avg_launch_ms'], d['ms_per_step'], d['roofline']['kernel_share_of_step'], d['acc_rate'])
PY
tail -3 gpurun_out/bench.err
B="python bench.py --steps 1 --warmup 1 --sites 2000000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"mh_update|csr_rebuild" -s 24 -c 2 -o gpurun_out/prof_mh $B > gpurun_out/ncu2.log 2>&1
