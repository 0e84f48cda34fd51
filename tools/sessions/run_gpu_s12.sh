# GPU-box session 12: update kernel with 64- and 32-thread CTAs (8 / 16 CTAs per SM) against 128 x 4
mkdir -p gpurun_out
V=$PWD/epievo_b200/variants; D=$PWD/epievo_b200/libepievo_b200.so
for v in t256 t512; do
  EPIEVO_B200_LIB=$V/lib_$v.so timeout 300 python -m pytest tests/test_gpu_chain_golden.py tests/test_gpu_sampler.py -m gpu -q -x > gpurun_out/r02_pytest_s12_$v.log 2>&1; echo "$v pytest rc=$?"; tail -1 gpurun_out/r02_pytest_s12_$v.log
done
run() { # label, lib, workload
  EPIEVO_B200_LIB=$2 timeout 300 python bench.py --workload $3 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s12_$1_$3.json 2> gpurun_out/bench_s12_$1_$3.err
  python -c "import json; d=json.loads([l for l in open('gpurun_out/bench_s12_$1_$3.json') if l.startswith('{')][-1]); print('$1 $3 value %.4g ms/step %.1f upd %.3f reb %.3f jd %.3f'%(d['value'], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['rebuild_kernel_avg_ms'], d['suffstats_pass']['ms']))"
}
for w in bal8 testtree bal32; do run default $D $w; run t256 $V/lib_t256.so $w; run t512 $V/lib_t512.so $w; done
