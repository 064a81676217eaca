# GPU-box recipe: parity, bench, then ncu launch list + full capture of the dominant kernel (run under gpurun)
python -m pytest tests/test_parity_gpu.py tests/test_scale_gpu.py -m gpu -q -x > gpurun_out/pytest_var.log 2>&1; tail -3 gpurun_out/pytest_var.log
python bench.py --points 200000000 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_c3_2e8.json 2> gpurun_out/bench.err; python -c "
import json; d=json.load(open('gpurun_out/bench_c3_2e8.json')); print(d['value'], d['roofline']['frac'], d['roofline']['stages_ms_per_step'])"
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_eval3d -s 6 -c 2 -o gpurun_out/prof_eval $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
