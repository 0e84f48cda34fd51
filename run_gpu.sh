# GPU-box session (round 2, session 5): no-alloc entry point, exact J/D over shards, -b over seeds, colour sets,
# drop-in CLI timing; A/B + profile of the mix update routine (default is the flat routine again)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_noalloc_alone.log
tail -4 gpurun_out/r02_pytest_noalloc_alone.log
timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_mcem.py tests/test_gpu_colour_sets.py tests/test_gpu_chain_golden.py tests/test_gpu_sharded.py tests/test_gpu_parity.py tests/test_gpu_compact.py tests/test_gpu_dropin_cli.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_pytest_s5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s5.log
tail -25 gpurun_out/r02_pytest_s5.log
timeout 900 python tools/cli_bench.py > gpurun_out/r02_cli_bench.json 2> gpurun_out/r02_cli_bench.err; tail -5 gpurun_out/r02_cli_bench.err; head -c 1500 gpurun_out/r02_cli_bench.json
for v in mix mixlocal; do
  EPIEVO_B200_LIB=$PWD/epievo_b200/variants/libepievo_$v.so timeout 300 python bench.py --workload bal8 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s5_${v}_bal8.json 2> gpurun_out/bench_s5_${v}_bal8.err
  python -c "import json; d=json.load(open('gpurun_out/bench_s5_${v}_bal8.json')); print('$v', d['roofline']['avg_launch_ms'])"
done
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
EPIEVO_B200_LIB=$PWD/epievo_b200/variants/libepievo_mix.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mh_update" -s 6 -c 2 -o gpurun_out/prof_mix $B > gpurun_out/ncu_mix.log 2>&1
tail -2 gpurun_out/ncu_mix.log
