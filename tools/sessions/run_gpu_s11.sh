# GPU-box session 11: est_complete drop-in test, CLI timings incl. est_complete, the whole CLI test file
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dropin_cli.py -m gpu -q > gpurun_out/r02_pytest_s11.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_s11.log
tail -5 gpurun_out/r02_pytest_s11.log | cut -c1-300
timeout 900 python tools/cli_bench.py > gpurun_out/r02_cli_bench.json 2> gpurun_out/r02_cli_bench.err; echo "cli_bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02_cli_bench.json'))
for k,v in d.items():
    if isinstance(v, dict): print(k, {a:(b.get('wall_s') if isinstance(b,dict) else b) for a,b in v.items()})
"
