# GPU-box session (round 2, final): the whole GPU suite, smoke(), the bench records of every workload, the
# reference arm, the ncu launch list + full captures of the three kernels of a step, a compute-sanitizer attempt
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_final.log
tail -4 gpurun_out/r02_pytest_gpu_final.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke_final.log 2>&1; tail -2 gpurun_out/r02_smoke_final.log
# 1. the driver's default command, then the other workloads (all legs: value, e2e, mcem, cpu_baseline)
timeout 600 python bench.py > gpurun_out/r02_bench_final_bal8.json 2> gpurun_out/r02_bench_final_bal8.err
python tools/show_bench.py gpurun_out/r02_bench_final_bal8.json
for w in testtree bal32 pairwise; do
  timeout 600 python bench.py --workload $w > gpurun_out/r02_bench_final_$w.json 2> gpurun_out/r02_bench_final_$w.err
  python tools/show_bench.py gpurun_out/r02_bench_final_$w.json
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_final_reference.json 2> gpurun_out/r02_bench_final_reference.err; tail -1 gpurun_out/r02_bench_final_reference.json | cut -c1-400
# 2. ncu: launch list of one warm-up + one timed step of the default command, full captures of the three kernels
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_final.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/ncu_list.log 2>&1; tail -1 gpurun_out/ncu_list.log | cut -c1-200
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"mh_update" -s 6 -c 1 -o gpurun_out/r02_prof_final_mh_update $B > gpurun_out/ncu_upd.log 2>&1; tail -1 gpurun_out/ncu_upd.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"csr_rebuild" -s 4 -c 1 -o gpurun_out/r02_prof_final_csr_rebuild $B > gpurun_out/ncu_reb.log 2>&1; tail -1 gpurun_out/ncu_reb.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"suffstats_tile" -s 1 -c 1 -o gpurun_out/r02_prof_final_suffstats_tile $B > gpurun_out/ncu_jd.log 2>&1; tail -1 gpurun_out/ncu_jd.log
# 3. compute-sanitizer (closed on this pool in every earlier session; the attempt and its message are the record)
timeout 300 compute-sanitizer --tool memcheck python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_sanitizer_memcheck.txt 2>&1; echo "rc=$?" >> gpurun_out/r02_sanitizer_memcheck.txt; tail -5 gpurun_out/r02_sanitizer_memcheck.txt | cut -c1-200
