# round 2, GPU call AD (2 GPUs): the default bench line launched the way the driver launches it for N = 2, both arms
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2ad_bench_2gpu.json 2> gpurun_out/r2ad_bench_2gpu.err
echo "rc $?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2ad_bench_2gpu.json'))
print('C3 2 GPUs', d['value'], d['n_gpus'], d['ms_per_step'], d['scaling'], (d.get('e2e') or {}).get('value'), d['clocks']['sm_mhz'], d['clocks']['reasons'])
for k,v in (d.get('workloads') or {}).items(): print(k, v.get('value'), (v.get('roofline') or {}).get('frac'))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2ad_bench_ref_2gpu.json 2> gpurun_out/r2ad_bench_ref_2gpu.err
echo "rc $?"; tail -c 300 gpurun_out/r2ad_bench_ref_2gpu.json
