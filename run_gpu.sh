# GPU-box session (round 2, session 6): whole GPU suite on the current default (flat routine), recovery of capacity
# overflow, drop-in CLI timing with the link-time writer, A/B + profile of the mix routine
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_s6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu_s6.log
tail -25 gpurun_out/r02_pytest_gpu_s6.log | cut -c1-250
timeout 900 python tools/cli_bench.py > gpurun_out/r02_cli_bench.json 2> gpurun_out/r02_cli_bench.err; tail -3 gpurun_out/r02_cli_bench.err
for v in mix mixlocal; do
  EPIEVO_B200_LIB=$PWD/epievo_b200/variants/libepievo_$v.so timeout 300 python bench.py --workload bal8 --no-cpu-baseline --no-mcem --no-e2e > gpurun_out/bench_s6_${v}_bal8.json 2> gpurun_out/bench_s6_${v}_bal8.err
  python -c "import json; d=json.load(open('gpurun_out/bench_s6_${v}_bal8.json')); print('$v', d['roofline']['avg_launch_ms'])"
done
timeout 300 python bench.py --workload bal8 > gpurun_out/bench_s6_bal8.json 2> gpurun_out/bench_s6_bal8.err; python -c "import json; d=json.load(open('gpurun_out/bench_s6_bal8.json')); print('default', d['value'], d['e2e']['value'], d['roofline']['avg_launch_ms'], d['mcem_iter']['ms'])"
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
EPIEVO_B200_LIB=$PWD/epievo_b200/variants/libepievo_mix.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mh_update" -s 6 -c 2 -o gpurun_out/prof_mix $B > gpurun_out/ncu_mix.log 2>&1
tail -2 gpurun_out/ncu_mix.log
