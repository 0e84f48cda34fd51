mkdir -p gpurun_out
python tools/indep_bench.py bal8 12500000 > gpurun_out/indep_bench.json 2> gpurun_out/indep_bench.err; tail -3 gpurun_out/indep_bench.err; cat gpurun_out/indep_bench.json
