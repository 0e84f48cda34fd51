# GPU-box session (round 2, session 7): the inactive-lane fix + restored deferral; A/B of two small variants
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sampler.py tests/test_gpu_chain_golden.py tests/test_gpu_multi.py tests/test_gpu_sharded.py tests/test_gpu_colour_sets.py tests/test_gpu_parity.py tests/test_gpu_compact.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_pytest_s7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s7.log
tail -15 gpurun_out/r02_pytest_s7.log | cut -c1-220
run() { # label, lib, workload
  EPIEVO_B200_LIB=$2 timeout 300 python bench.py --workload $3 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s7_$1_$3.json 2> gpurun_out/bench_s7_$1_$3.err
  python -c "import json; d=json.load(open('gpurun_out/bench_s7_$1_$3.json')); print('$1 $3 value %.4g upd %.3f reb %.3f jd %.3f'%(d['value'], d['roofline']['avg_launch_ms'], d['roofline']['rebuild_kernel_avg_ms'], d['suffstats_pass']['ms']))"
}
D=$PWD/epievo_b200/libepievo_b200.so; V=$PWD/epievo_b200/variants
run default $D bal8; run jbase $V/libepievo_jbase.so bal8; run nosplit $V/libepievo_nosplit.so bal8
run default $D testtree; run jbase $V/libepievo_jbase.so testtree; run nosplit $V/libepievo_nosplit.so testtree
run default $D bal32; run nosplit $V/libepievo_nosplit.so bal32
run default $D pairwise
