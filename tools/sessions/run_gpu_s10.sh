# GPU-box session 10: the late capacity watch on long rows (new test), halo capacity sized from the branch lengths,
# and where the end-to-end leg of the small config goes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multi.py -m gpu -q > gpurun_out/r02_pytest_s10.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s10.log
tail -5 gpurun_out/r02_pytest_s10.log | cut -c1-300
timeout 300 python tools/e2e_probe.py --workload pairwise > gpurun_out/r02_e2e_probe_pairwise.txt 2>&1; cat gpurun_out/r02_e2e_probe_pairwise.txt | cut -c1-400
timeout 300 python tools/e2e_probe.py --workload bal8 --iters 3 > gpurun_out/r02_e2e_probe_bal8.txt 2>&1; cat gpurun_out/r02_e2e_probe_bal8.txt | cut -c1-400
