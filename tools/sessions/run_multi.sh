# Multi-GPU session: bash run_multi.sh N   (gpurun --gpus N; charged N x the box time, so every step is bounded)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r02_topo_n$N.txt 2>&1
# 1. N ranks over real NCCL == one GPU, bit for bit (overlapped + serial protocol, -b loop)
if [ "$2" != "bench-only" ]; then
timeout 400 $TR --master-port 29511 tools/dist_check.py > gpurun_out/r02_dist_check_n$N.txt 2>&1; echo "rc=$?" >> gpurun_out/r02_dist_check_n$N.txt
grep -E "BIT-EXACT|MISMATCH|rc=" gpurun_out/r02_dist_check_n$N.txt
fi
# 2. weak scaling, headline shard per GPU, overlapped exchange (and the serial one for comparison at N=2, 8)
timeout 600 $TR --master-port 29512 bench.py --gpus $N --no-cpu-baseline > gpurun_out/r02_bench_bal8_weak_n$N.json 2> gpurun_out/r02_bench_bal8_weak_n$N.err
python tools/show_bench.py gpurun_out/r02_bench_bal8_weak_n$N.json
if [ "$2" = "serial" ]; then
  timeout 600 $TR --master-port 29513 bench.py --gpus $N --no-cpu-baseline --no-e2e --no-mcem --serial-exchange > gpurun_out/r02_bench_bal8_weak_serial_n$N.json 2> gpurun_out/r02_bench_bal8_weak_serial_n$N.err
  python tools/show_bench.py gpurun_out/r02_bench_bal8_weak_serial_n$N.json
fi
# 3. config 4 as written: 3e7 sites in total, strong scaling
timeout 600 $TR --master-port 29514 bench.py --gpus $N --workload testtree --scaling strong --sites 30000000 --no-cpu-baseline > gpurun_out/r02_bench_c4_strong_n$N.json 2> gpurun_out/r02_bench_c4_strong_n$N.err
python tools/show_bench.py gpurun_out/r02_bench_c4_strong_n$N.json
if [ "$N" = "8" ] && [ "$2" != "bench-only" ]; then
  # 4. config 5 at its stated shape: 32-leaf tree, 1e8 sites, -B 100 -L 20, -b, 8 GPUs (histories generated on the device)
  timeout 900 $TR --master-port 29515 bench.py --gpus 8 --workload bal32 --scaling strong --sites 100000000 --burn-in 20 --batch 100 --steps 1 --warmup 1 --optimize-branches --gen device --no-cpu-baseline > gpurun_out/r02_bench_c5_n8.json 2> gpurun_out/r02_bench_c5_n8.err
  python tools/show_bench.py gpurun_out/r02_bench_c5_n8.json
  tail -3 gpurun_out/r02_bench_c5_n8.err
fi
if [ "$N" = "2" ]; then
  # the drop-in binary on 2 GPUs (EPIEVO_B200_GPUS), plain C++ over the C-ABI
  timeout 120 tests/cpp/_build/test_multi 2 200000 > gpurun_out/r02_cpp_test_multi_n2.txt 2>&1; cat gpurun_out/r02_cpp_test_multi_n2.txt
  timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/r02_pytest_multi_n2.log 2>&1; tail -3 gpurun_out/r02_pytest_multi_n2.log
fi
