# round 2, multi-GPU call (N = number of GPUs of the box, passed as $1): torchrun bench, one-call-all-GPUs host path, sharded entry
N=${1:-2}
timeout 600 python -m pytest tests/test_scale_gpu.py -m gpu -q -k "sharded or C3" > gpurun_out/r2g_pytest_${N}gpu.log 2>&1; tail -3 gpurun_out/r2g_pytest_${N}gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2g_bench_${N}gpu.json 2> gpurun_out/r2g_bench_${N}gpu.err; cut -c1-400 gpurun_out/r2g_bench_${N}gpu.json; tail -2 gpurun_out/r2g_bench_${N}gpu.err
AB_POINTS=400000000 timeout 900 python scripts/host_path_ab.py > gpurun_out/r2g_host_ab_${N}gpu.txt 2>&1; cat gpurun_out/r2g_host_ab_${N}gpu.txt
timeout 900 python bench.py --steps 5 --no-extra > gpurun_out/r2g_bench_1proc_${N}visible.json 2> gpurun_out/r2g_bench_1proc.err; python - <<PY
import json
d=json.load(open('gpurun_out/r2g_bench_1proc_${N}visible.json')); print('single process, $N GPUs visible: value %.3e e2e %.3e pageable %.3e all_gpus %s' % (d['value'], d['e2e']['value'], d['e2e']['pageable']['value'], d['e2e']['all_gpus']))
PY
