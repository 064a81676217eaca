# round 2, GPU call AE: ncu of the cooperative sort launch and the 2D evaluation at C1's own size (1e6 points)
mkdir -p gpurun_out
CMD="python bench.py --workload C1 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2ae_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_sort_fused -s 3 -c 1 -f -o gpurun_out/r2ae_prof_c1_sort $CMD > gpurun_out/r2ae_ncu1.log 2>&1
tail -1 gpurun_out/r2ae_ncu1.log
$CMD > gpurun_out/r2ae_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_eval2d -s 3 -c 1 -f -o gpurun_out/r2ae_prof_c1_eval $CMD > gpurun_out/r2ae_ncu2.log 2>&1
tail -1 gpurun_out/r2ae_ncu2.log
