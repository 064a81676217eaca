run() { # name, env..., args
  name=$1; shift
  env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${name}.json 2> gpurun_out/var_${name}.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${name}.json')); r=d['roofline']; print('${name} value=%.3e ms/step=%.2f frac=%.3f'%(d['value'], d['ms_per_step'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()}, d['clocks']['sm_mhz'])" || tail -3 gpurun_out/var_${name}.err
}
python -m pytest tests/test_parity_gpu.py tests/test_scale_gpu.py -m gpu -q -x > gpurun_out/pytest_var.log 2>&1; tail -2 gpurun_out/pytest_var.log
ARGS="--points 200000000"
run coal X=1
run nocoal BAOBZI_B200_SORT_COAL=0
