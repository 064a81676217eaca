timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 600 python bench.py --workload C1 --steps 20 --warmup 5 > gpurun_out/wl_C1.json 2> gpurun_out/wl_C1.err; cut -c1-200 gpurun_out/wl_C1.json
