# round 2, GPU call AG (8 GPUs): the default bench line launched the way the driver launches it for N = 8
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2ag_bench_8gpu.json 2> gpurun_out/r2ag_bench_8gpu.err
echo "rc $?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2ag_bench_8gpu.json'))
print('C3 8 GPUs', d['value'], d['n_gpus'], d['ms_per_step'], d['scaling'], (d.get('e2e') or {}).get('value'), d['clocks']['sm_mhz'], d['clocks']['reasons'])
for k,v in (d.get('workloads') or {}).items(): print(k, v.get('value'))
PY
