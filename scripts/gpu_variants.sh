# GPU-box experiment: K1/K3 privatised vs global atomics
python -m pytest tests/test_parity_gpu.py tests/test_scale_gpu.py -m gpu -q -x > gpurun_out/pytest_var.log 2>&1; tail -3 gpurun_out/pytest_var.log
run() { # name, env..., args
  name=$1; shift
  env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${name}.json 2> gpurun_out/var.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${name}.json')); r=d['roofline']; print('${name} value=%.3e frac=%.3f'%(d['value'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()})"
}
ARGS="--points 200000000"
run c3_priv BAOBZI_B200_PRIVATE=1
run c3_glob BAOBZI_B200_PRIVATE=0
ARGS="--points 100000000 --workload C5"
run c5_priv BAOBZI_B200_PRIVATE=1
run c5_glob BAOBZI_B200_PRIVATE=0
