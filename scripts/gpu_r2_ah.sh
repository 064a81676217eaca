# round 2, GPU call AH: ncu --set full of the headline evaluation kernel (K4) and of the two-leaf 1D kernel on the last build
mkdir -p gpurun_out
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2ah_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_eval3d_kernel -s 6 -c 1 -f -o gpurun_out/r2ah_prof_k4 $CMD > gpurun_out/r2ah_ncu1.log 2>&1
tail -1 gpurun_out/r2ah_ncu1.log
CMD2="python bench.py --workload C2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD2 > gpurun_out/r2ah_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:two_leaf -s 3 -c 1 -f -o gpurun_out/r2ah_prof_twoleaf $CMD2 > gpurun_out/r2ah_ncu2.log 2>&1
tail -1 gpurun_out/r2ah_ncu2.log
