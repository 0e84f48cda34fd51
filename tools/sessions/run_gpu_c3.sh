# GPU-box session: config 3 at its stated shape (1e8 sites, test tree, one GPU) and the per-site slowdown of the update
# kernel at 1e8 sites: the same kernel profiled at 3e7 and at 1e8 sites
mkdir -p gpurun_out
timeout 600 python bench.py --workload testtree --sites 100000000 --gen device --no-cpu-baseline --no-mcem > gpurun_out/r02_bench_c3_1e8.json 2> gpurun_out/r02_bench_c3_1e8.err
python -c "import json; d=json.load(open('gpurun_out/r02_bench_c3_1e8.json')); print('C3 1e8: jd', d['suffstats_pass'], 'upd', d['roofline']['avg_launch_ms'], 'reb', d['roofline']['rebuild_kernel_avg_ms'], d['value'])"
timeout 300 python bench.py --workload testtree --gen device --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/r02_bench_c4_3e7_devgen.json 2> gpurun_out/r02_bench_c4_3e7_devgen.err
python -c "import json; d=json.load(open('gpurun_out/r02_bench_c4_3e7_devgen.json')); print('3e7 devgen: upd', d['roofline']['avg_launch_ms'], 'reb', d['roofline']['rebuild_kernel_avg_ms'], d['value'])"
B="python bench.py --workload testtree --gen device --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
timeout 600 ncu --set full --clock-control none -k regex:"mh_update" -s 6 -c 1 -o gpurun_out/prof_tt_3e7 $B > gpurun_out/ncu_tt_3e7.log 2>&1; tail -1 gpurun_out/ncu_tt_3e7.log
timeout 900 ncu --set full --clock-control none -k regex:"mh_update" -s 6 -c 1 -o gpurun_out/prof_tt_1e8 $B --sites 100000000 > gpurun_out/ncu_tt_1e8.log 2>&1; tail -1 gpurun_out/ncu_tt_1e8.log
