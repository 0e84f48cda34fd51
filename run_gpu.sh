# GPU-box session (round 2, session 4): mix update kernel (default) vs flat; multi-GPU handle with exact J/D
mkdir -p gpurun_out
FLAT=$PWD/epievo_b200/variants/libepievo_flat.so
EPIEVO_B200_LIB=$FLAT timeout 600 python -m pytest tests/test_gpu_chain_golden.py tests/test_gpu_multi.py tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/r02_pytest_s4_flat.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s4_flat.log
tail -6 gpurun_out/r02_pytest_s4_flat.log
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_chain_golden.py > gpurun_out/r02_pytest_s4_mix.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s4_mix.log
tail -30 gpurun_out/r02_pytest_s4_mix.log
for w in bal8 testtree bal32 pairwise; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s4_$w.json 2> gpurun_out/bench_s4_$w.err; tail -c 700 gpurun_out/bench_s4_$w.json | head -c 400; echo
done
timeout 900 python tools/stats_probe.py --arm gpu --R 2048 --burn 300 > gpurun_out/r02_stats_probe_mix_bal8_R2048_burn300.txt 2>&1
tail -6 gpurun_out/r02_stats_probe_mix_bal8_R2048_burn300.txt
