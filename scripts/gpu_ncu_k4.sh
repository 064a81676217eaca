# ncu: launch list, then one full capture (with source counters) of the dominant kernel
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile ${BENCH_ARGS}"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-bin_eval3d} -s 6 -c 1 -f -o gpurun_out/prof_eval $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
