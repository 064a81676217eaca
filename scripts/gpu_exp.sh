# scratch experiment: 2D evaluation, streaming result stores, chunk sweep
for c in 2097152 3145728 4194304; do
BAOBZI_B200_CHUNK=$c timeout 300 python bench.py --workload C1 --points 100000000 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/c1_ring_$c.json 2> gpurun_out/c1_ring.err
python -c "
import json; d=json.load(open('gpurun_out/c1_ring_$c.json')); r=d['roofline']; print('C1 stcs chunk $c value=%.3e ms=%.3f'%(d['value'], d['ms_per_step']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/c1_ring.err
done
