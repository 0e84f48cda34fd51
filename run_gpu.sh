# GPU-box session (round 2, session 3): multi-GPU handle + overlapped protocol on one GPU; ncu of the lockstep variant
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_sharded.py tests/test_gpu_chain_golden.py tests/test_gpu_compact.py tests/test_gpu_indep.py tests/test_gpu_dropin_cli.py tests/test_gpu_mcem.py -m gpu -x -q > gpurun_out/r02_pytest_s3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s3.log
tail -25 gpurun_out/r02_pytest_s3.log
timeout 300 python bench.py --workload bal8 --no-cpu-baseline --no-mcem > gpurun_out/bench_s3_bal8.json 2> gpurun_out/bench_s3_bal8.err; tail -c 300 gpurun_out/bench_s3_bal8.json; echo
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
EPIEVO_B200_LIB=$PWD/epievo_b200/variants/libepievo_lock.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mh_update" -s 6 -c 2 -o gpurun_out/prof_lock $B > gpurun_out/ncu_lock.log 2>&1
tail -3 gpurun_out/ncu_lock.log
