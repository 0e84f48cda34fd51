# GPU-box session 14: why do larger CTAs run the update kernel faster?  ncu on the test tree, 128 x 4 vs 512 x 1
mkdir -p gpurun_out
B="python bench.py --workload testtree --steps 1 --warmup 1 --burn-in 1 --batch 1 --no-e2e --no-cpu-baseline --no-mcem"
timeout 400 ncu --set full --clock-control none -k regex:"mh_update" -s 6 -c 1 -o gpurun_out/prof_tt_t128 $B > gpurun_out/ncu_t128.log 2>&1; tail -1 gpurun_out/ncu_t128.log
EPIEVO_B200_LIB=$PWD/epievo_b200/variants/lib_t512.so timeout 400 ncu --set full --clock-control none -k regex:"mh_update" -s 6 -c 1 -o gpurun_out/prof_tt_t512 $B > gpurun_out/ncu_t512.log 2>&1; tail -1 gpurun_out/ncu_t512.log
