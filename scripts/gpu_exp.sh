# scratch experiment: device hooks + full GPU suite + C2 line
timeout 900 python -m pytest tests/test_device_hook.py -x -q -m gpu 2>&1 | tail -8
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --workload C2 --steps 5 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; tail -c 600 gpurun_out/bench_c2.json
