# round 2, GPU call A: the whole GPU suite (new: reference consumers, C4 / 65 K-leaf trees, alignment), smoke, both
# bench arms (default invocation = headline C3 + the other four configurations), memcheck
timeout 2400 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r2a_pytest.log 2>&1; tail -25 gpurun_out/r2a_pytest.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/r2a_smoke.log 2>&1; tail -1 gpurun_out/r2a_smoke.log
bash scripts/gpu_r2_b.sh 2>&1 | tee gpurun_out/r2b_sweep.txt
timeout 1200 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; cut -c1-600 gpurun_out/r2a_bench.json; tail -3 gpurun_out/r2a_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; cut -c1-300 gpurun_out/r2a_bench_ref.json
timeout 1500 compute-sanitizer --tool memcheck python scripts/sanitize_paths.py > gpurun_out/r2a_memcheck.txt 2>&1; tail -4 gpurun_out/r2a_memcheck.txt
nvidia-smi --query-gpu=name,memory.used,clocks.sm --format=csv
