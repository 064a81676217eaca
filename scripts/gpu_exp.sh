for o in 6 10 16; do
for m in binned direct; do
timeout 600 python bench.py --workload C1 --order $o --mode $m --points 100000000 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/x.json 2>gpurun_out/x.err; python -c "
import json; d=json.load(open('gpurun_out/x.json')); r=d['roofline']; print('C1 fn order $o $m value=%.4e ms=%.3f'%(d['value'], d['ms_per_step']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/x.err
done; done
