mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"mh_update|csr_rebuild|suffstats_kernel" -s 13 -c 13 -o gpurun_out/prof_v6 $B > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
