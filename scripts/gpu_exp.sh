timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -2
timeout 600 python bench.py --workload C3 --order 6 --points 200000000 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/c3_o6.json 2>gpurun_out/c3_o6.err; python -c "
import json; d=json.load(open('gpurun_out/c3_o6.json')); r=d['roofline']; print('3D order 6 value=%.4e ms=%.3f frac=%.3f'%(d['value'], d['ms_per_step'], r['frac']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, r['points_per_launch'])" || tail -3 gpurun_out/c3_o6.err
