N=${NG:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload C5 --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err
cut -c1-300 gpurun_out/bench_c5_n$N.json; tail -2 gpurun_out/bench_c5_n$N.err
