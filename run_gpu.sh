mkdir -p gpurun_out
python -m pytest tests/test_gpu_sharded.py tests/test_gpu_compact.py tests/test_gpu_sampler.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err || tail -5 gpurun_out/bench_n2.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
print('N=2', '%.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'ms/step %.1f'%d['ms_per_step'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'])
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/dist_check.py > gpurun_out/dist_check_n2.txt 2>&1; tail -3 gpurun_out/dist_check_n2.txt
