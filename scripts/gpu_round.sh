# GPU-box recipe for the judged artefacts: tests, smoke, both bench arms, launch list, full ncu captures
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; cut -c1-400 gpurun_out/bench_default.json; tail -2 gpurun_out/bench_default.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-300 gpurun_out/bench_ref.json
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile"
$CMD > gpurun_out/plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bin_eval3d -s 6 -c 1 -f -o gpurun_out/prof_k4 $CMD > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bin_sort_fused -s 6 -c 2 -f -o gpurun_out/prof_sort $CMD > gpurun_out/ncu3.log 2>&1
tail -1 gpurun_out/ncu3.log
