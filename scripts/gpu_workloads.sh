# GPU-box recipe: every BASELINE.json configuration once (device-resident, no CPU baseline)
for w in C1 C2 C4 C5; do
  python bench.py --workload $w --steps 3 --warmup 3 --no-cpu ${EXTRA_ARGS} > gpurun_out/wl_$w.json 2> gpurun_out/wl_$w.err
  python -c "
import json; d=json.load(open('gpurun_out/wl_$w.json')); r=d['roofline']; print('$w value=%.3e e2e=%.3e bound=%s frac=%.3f'%(d['value'], (d.get('e2e') or {}).get('value',0), r['bound'], r['frac']), r.get('kernel'), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()}, 'hbm_frac=%.3f'%r.get('hbm_frac',0), d['config'].get('fit_seconds'))" || tail -3 gpurun_out/wl_$w.err
done
