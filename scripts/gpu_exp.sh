timeout 900 python -m pytest tests/test_scale_gpu.py -x -q -m gpu 2>&1 | tail -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
cut -c1-330 gpurun_out/bench_n2.json; tail -2 gpurun_out/bench_n2.err
