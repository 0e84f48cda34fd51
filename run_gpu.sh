# One GPU-box session: parity tests, the bench records and the ncu evidence behind profiles/.
# usage: gpurun --timeout 1800 -- 'bash run_gpu.sh'   (outputs land in gpurun_out/)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -3 gpurun_out/pytest.log
for w in bal8 testtree bal32 pairwise; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 200 gpurun_out/bench_$w.json
done
# launch list of two timed steps (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-mcem > gpurun_out/ncu_launch.log 2>&1
# full capture of one E-step's kernels (burn-in 1, batch 1: 6 update + 6 overflow + 6 rebuild + the J/D pass)
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mh_update|csr_rebuild|suffstats_tile|suffstats_kernel" \
  -s 21 -c 21 -o gpurun_out/prof $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
