mkdir -p gpurun_out
for n in 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err || tail -5 gpurun_out/bench_n$n.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1])
print('N=$n', '%.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'ms/step %.1f'%d['ms_per_step'], 'upd_ms=%.3f'%d['roofline']['avg_launch_ms'])
PY
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 tools/dist_check.py > gpurun_out/dist_check_n8.txt 2>&1; tail -3 gpurun_out/dist_check_n8.txt
