run() { # name, env..., args
  name=$1; shift
  env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${name}.json 2> gpurun_out/var_${name}.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${name}.json')); r=d['roofline']; print('${name} value=%.3e frac=%.3f'%(d['value'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()}, d['clocks']['sm_mhz'])" || tail -3 gpurun_out/var_${name}.err
}
ARGS="--points 200000000"
for v in 4,8,0,1 91,8,0,1 92,8,0,1 91,12,0,0 92,12,0,0; do run k4_${v//,/_} BAOBZI_B200_K4=$v; done
