# GPU-box experiment: K4 variants (BAOBZI_B200_K4="P,warps,tzfly,tykeep") and sort CTAs per SM
run() { # name, env..., args
  name=$1; shift
  env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${name}.json 2> gpurun_out/var_${name}.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${name}.json')); r=d['roofline']; print('${name} value=%.3e frac=%.3f'%(d['value'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()}, d['clocks']['sm_mhz'])" || tail -3 gpurun_out/var_${name}.err
}
python -m pytest tests/test_parity_gpu.py -m gpu -q -x > gpurun_out/pytest_var.log 2>&1; tail -3 gpurun_out/pytest_var.log
ARGS="--points 200000000"
run default X=1
for v in 4,8,0,0 3,12,0,0 3,12,0,1 4,12,1,0 2,16,0,0 2,16,0,1 3,8,0,1; do run k4_${v//,/_} BAOBZI_B200_K4=$v; done
for c in 1 2 3; do run ctas$c BAOBZI_B200_SORT_CTAS=$c; done
run legacy_sort BAOBZI_B200_FUSED_SORT=0
