# GPU-box session: the bulk-staged rebuild kernel (EPIEVO_B200_REBUILD=bulk[:BC]) against the default one --
# chains must be identical (chain goldens, sampler / parity / multi tests), then per-pass times per workload
mkdir -p gpurun_out
EPIEVO_B200_REBUILD=bulk:4 timeout 900 python -m pytest tests/test_gpu_chain_golden.py tests/test_gpu_sampler.py tests/test_gpu_parity.py tests/test_gpu_compact.py tests/test_gpu_multi.py tests/test_gpu_sharded.py -m gpu -q -x > gpurun_out/r02_pytest_bulk.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_bulk.log
tail -6 gpurun_out/r02_pytest_bulk.log | cut -c1-200
run() { # label, mode, workload
  EPIEVO_B200_REBUILD=$2 timeout 300 python bench.py --workload $3 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_rb_$1_$3.json 2> gpurun_out/bench_rb_$1_$3.err
  python -c "import json; d=json.loads([l for l in open('gpurun_out/bench_rb_$1_$3.json') if l.startswith('{')][-1]); print('$1 $3 value %.4g upd %.3f reb %.3f jd %.3f'%(d['value'], d['roofline']['avg_launch_ms'], d['roofline']['rebuild_kernel_avg_ms'], d['suffstats_pass']['ms']))"
}
run legacy legacy bal8; run bulk2 bulk:2 bal8; run bulk4 bulk:4 bal8; run bulk7 bulk:7 bal8
run legacy legacy testtree; run bulk2 bulk:2 testtree; run bulk4 bulk:4 testtree
run legacy legacy bal32; run bulk4 bulk:4 bal32; run bulk8 bulk:8 bal32
