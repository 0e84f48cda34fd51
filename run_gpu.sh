mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_bal8.json 2> gpurun_out/bench_bal8.err; tail -c 600 gpurun_out/bench_bal8.json
for w in testtree bal32 pairwise; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 300 gpurun_out/bench_$w.json
done
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v7.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-mcem > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mh_update|csr_rebuild|suffstats_tile|suffstats_kernel" -s 15 -c 15 -o gpurun_out/prof_v7 $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
