# GPU-box experiment: CTAs per SM of the privatised K1/K3
run() { # name, env..., args
  name=$1; shift
  env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${name}.json 2> gpurun_out/var.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${name}.json')); r=d['roofline']; print('${name} value=%.3e frac=%.3f'%(d['value'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()}, d['clocks'])"
}
ARGS="--points 200000000"
for c in 1 2 4 8; do run c3_ctas$c BAOBZI_B200_SORT_CTAS=$c; done
run c3_glob BAOBZI_B200_PRIVATE=0
