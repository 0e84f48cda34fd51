# GPU-box session (round 2, final build with one large CTA per SM): tests that run 64-node trees, bench records of
# every workload, ncu launch list + full capture of the update kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_chain_golden.py tests/test_gpu_sampler.py tests/test_gpu_sharded.py tests/test_gpu_multi.py tests/test_gpu_colour_sets.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02_pytest_final2.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r02_pytest_final2.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke_final2.log 2>&1; tail -1 gpurun_out/r02_smoke_final2.log
timeout 600 python bench.py > gpurun_out/r02_bench_final2_bal8.json 2> gpurun_out/r02_bench_final2_bal8.err
python tools/show_bench.py gpurun_out/r02_bench_final2_bal8.json
for w in testtree bal32 pairwise; do
  timeout 600 python bench.py --workload $w > gpurun_out/r02_bench_final2_$w.json 2> gpurun_out/r02_bench_final2_$w.err
  python tools/show_bench.py gpurun_out/r02_bench_final2_$w.json
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_final2.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/ncu_list.log 2>&1; tail -1 gpurun_out/ncu_list.log | cut -c1-200
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"mh_update" -s 6 -c 1 -o gpurun_out/r02_prof_final2_mh_update $B > gpurun_out/ncu_upd.log 2>&1; tail -1 gpurun_out/ncu_upd.log
