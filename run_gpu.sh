mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -3 gpurun_out/pytest.log
timeout 300 python bench.py --workload pairwise > gpurun_out/bench_pairwise.json 2> gpurun_out/bench_pairwise.err; tail -c 200 gpurun_out/bench_pairwise.json
timeout 300 python bench.py > gpurun_out/bench_bal8.json 2> gpurun_out/bench_bal8.err; tail -c 200 gpurun_out/bench_bal8.json
