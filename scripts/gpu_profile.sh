# GPU-box recipe: full bench line, then ncu launch list + full capture of the dominant kernel (run under gpurun)
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c3_full.json 2> gpurun_out/bench_c3_full.err; tail -c 1200 gpurun_out/bench_c3_full.json; tail -3 gpurun_out/bench_c3_full.err
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_eval3d -s 6 -c 2 -o gpurun_out/prof_eval $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
