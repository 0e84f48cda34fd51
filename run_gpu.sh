# One GPU-box session (round 2, session 1): the sampler-law evidence.
# usage: gpurun --timeout 1500 -- 'bash run_gpu.sh'   (outputs land in gpurun_out/)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_exact_posterior.py tests/test_gpu_sampler.py -m gpu -q -s > gpurun_out/r02_pytest_law.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_law.log
tail -5 gpurun_out/r02_pytest_law.log
for burn in 30 300; do
  timeout 600 python tools/stats_probe.py --arm gpu --R 512 --burn $burn > gpurun_out/r02_stats_probe_bal8_R512_burn$burn.txt 2>&1
  tail -3 gpurun_out/r02_stats_probe_bal8_R512_burn$burn.txt
done
# compute-sanitizer on the smoke path (memcheck, then racecheck): the colour classes' race freedom
for tool in memcheck racecheck; do
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python __graft_entry__.py --smoke > gpurun_out/r02_sanitizer_${tool}_smoke.txt 2>&1
  echo "sanitizer $tool rc=$?" >> gpurun_out/r02_sanitizer_${tool}_smoke.txt
  tail -4 gpurun_out/r02_sanitizer_${tool}_smoke.txt
done
