mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
B="python bench.py --workload bal32 --steps 1 --warmup 1 --sites 400000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"mh_update|suffstats_kernel" -s 20 -c 4 -o gpurun_out/prof_b32 $B > gpurun_out/ncu2.log 2>&1
B="python bench.py --steps 1 --warmup 1 --sites 2000000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"suffstats_kernel" -s 1 -c 1 -o gpurun_out/prof_ss $B > gpurun_out/ncu3.log 2>&1
