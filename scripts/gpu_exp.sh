timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/c3_agg.json 2>gpurun_out/c3_agg.err; python -c "
import json; d=json.load(open('gpurun_out/c3_agg.json')); r=d['roofline']; print('C3 value=%.4e ms=%.3f frac=%.3f'%(d['value'], d['ms_per_step'], r['frac']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()})"
timeout 600 python bench.py --workload C5 --steps 2 --warmup 2 --no-cpu --no-e2e > gpurun_out/c5_agg.json 2>gpurun_out/c5_agg.err; python -c "
import json; d=json.load(open('gpurun_out/c5_agg.json')); r=d['roofline']; print('C5 value=%.4e ms=%.3f frac=%.3f'%(d['value'], d['ms_per_step'], r['frac']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()})"
