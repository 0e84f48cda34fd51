mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sim.py -x -q > gpurun_out/pytest_sim.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sim.log
tail -15 gpurun_out/pytest_sim.log
timeout 300 python tools/sim_bench.py testtree 30000000 > gpurun_out/sim_testtree.json 2> gpurun_out/sim_testtree.err; cat gpurun_out/sim_testtree.json; tail -3 gpurun_out/sim_testtree.err
timeout 300 python tools/sim_bench.py bal8 12500000 > gpurun_out/sim_bal8.json 2> gpurun_out/sim_bal8.err; cat gpurun_out/sim_bal8.json; tail -3 gpurun_out/sim_bal8.err
