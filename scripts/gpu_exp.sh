for p in 32768 65536 131072 200000; do
for m in direct binned; do
timeout 300 python bench.py --workload C3 --points $p --mode $m --steps 20 --warmup 5 --no-cpu --no-e2e > gpurun_out/c3_x.json 2> gpurun_out/c3_x.err
python -c "
import json; d=json.load(open('gpurun_out/c3_x.json')); r=d['roofline']; print('C3 $m $p value=%.3e ms=%.4f'%(d['value'], d['ms_per_step']), {k:round(v,4) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/c3_x.err
done; done
for p in 16384 65536; do
for m in direct binned; do
timeout 300 python bench.py --workload C5 --points $p --mode $m --steps 20 --warmup 5 --no-cpu --no-e2e > gpurun_out/c3_x.json 2> gpurun_out/c3_x.err
python -c "
import json; d=json.load(open('gpurun_out/c3_x.json')); r=d['roofline']; print('C5 $m $p value=%.3e ms=%.4f'%(d['value'], d['ms_per_step']), {k:round(v,4) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/c3_x.err
done; done
