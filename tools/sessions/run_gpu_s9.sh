# GPU-box session 9: rebuild kernel with batched first-element loads (default) vs the previous write phase (rbold) and
# two register caps; lagged per-sweep housekeeping (chains must not change: chain goldens)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain_golden.py tests/test_gpu_sampler.py tests/test_gpu_parity.py tests/test_gpu_compact.py tests/test_gpu_multi.py tests/test_gpu_sharded.py tests/test_gpu_mcem.py tests/test_gpu_colour_sets.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02_pytest_s9.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s9.log
tail -5 gpurun_out/r02_pytest_s9.log | cut -c1-200
run() { # label, lib, workload
  EPIEVO_B200_LIB=$2 timeout 300 python bench.py --workload $3 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s9_$1_$3.json 2> gpurun_out/bench_s9_$1_$3.err
  python -c "import json; d=json.loads([l for l in open('gpurun_out/bench_s9_$1_$3.json') if l.startswith('{')][-1]); print('$1 $3 value %.4g ms/step %.1f upd %.3f reb %.3f jd %.3f'%(d['value'], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['rebuild_kernel_avg_ms'], d['suffstats_pass']['ms']))"
}
D=$PWD/epievo_b200/libepievo_b200.so; V=$PWD/epievo_b200/variants
for w in bal8 testtree bal32 pairwise; do
  run default $D $w; run rbold $V/lib_rbold.so $w; run rbmb4 $V/lib_rbmb4.so $w; run rbmb5 $V/lib_rbmb5.so $w
done
