# GPU-box session 15: one large CTA per SM for the update kernel's main launch (512 threads up to 16 nodes, 384 up to
# 64): the whole GPU suite, the four workloads, and 512 threads for the 64-node variant
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_s15.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r02_pytest_s15.log | cut -c1-200
run() { # label, lib, workload
  EPIEVO_B200_LIB=$2 timeout 300 python bench.py --workload $3 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s15_$1_$3.json 2> gpurun_out/bench_s15_$1_$3.err
  python -c "import json; d=json.loads([l for l in open('gpurun_out/bench_s15_$1_$3.json') if l.startswith('{')][-1]); print('$1 $3 value %.4g ms/step %.1f upd %.3f reb %.3f jd %.3f'%(d['value'], d['ms_per_step'], d['roofline']['avg_launch_ms'], d['roofline']['rebuild_kernel_avg_ms'], d['suffstats_pass']['ms']))"
}
D=$PWD/epievo_b200/libepievo_b200.so; V=$PWD/epievo_b200/variants
for w in bal8 testtree bal32 pairwise; do run default $D $w; done
run t64_512 $V/lib_t64_512.so bal32
