mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
nproc; free -g | head -2
python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench.log
B="python bench.py --steps 1 --warmup 1 --sites 2000000 --burn-in 2 --batch 2 --no-e2e --no-cpu-baseline"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mh_pass -s 12 -c 3 -o gpurun_out/prof_mh $B > gpurun_out/ncu2.log 2>&1
$B > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:suffstats_kernel -s 1 -c 1 -o gpurun_out/prof_ss $B > gpurun_out/ncu3.log 2>&1
echo done
