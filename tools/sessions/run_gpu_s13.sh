# GPU-box session 13: persistent main launch of the update kernel (EPIEVO_B200_PERSIST=0: one CTA per tile as before)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain_golden.py tests/test_gpu_sampler.py tests/test_gpu_colour_sets.py tests/test_gpu_multi.py tests/test_gpu_sharded.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02_pytest_s13.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r02_pytest_s13.log
run() { # persist, workload
  EPIEVO_B200_PERSIST=$1 timeout 300 python bench.py --workload $2 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s13_p$1_$2.json 2> gpurun_out/bench_s13_p$1_$2.err
  python -c "import json; d=json.loads([l for l in open('gpurun_out/bench_s13_p$1_$2.json') if l.startswith('{')][-1]); print('persist=$1 $2 value %.4g ms/step %.1f upd %.3f reb %.3f jd %.3f'%(d['value'], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['rebuild_kernel_avg_ms'], d['suffstats_pass']['ms']))"
}
for w in bal8 testtree bal32 pairwise; do for p in 0 1 2 4; do run $p $w; done; done
