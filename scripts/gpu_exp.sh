export BAOBZI_B200_BCHUNK=16777216 BAOBZI_B200_BUCKETS=4096
CMD="python bench.py --workload C4 --mode bucketed --points 67108864 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:eval_bucket -s 6 -c 1 -f -o gpurun_out/prof_perm $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
