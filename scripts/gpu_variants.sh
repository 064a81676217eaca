# GPU-box experiment: K4 variants on workload C3 (and C5)
python -m pytest tests/test_parity_gpu.py tests/test_scale_gpu.py -m gpu -q -x > gpurun_out/pytest_var.log 2>&1; tail -3 gpurun_out/pytest_var.log
run() { # name, env..., args
  name=$1; shift
  env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${name}.json 2> gpurun_out/var.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${name}.json')); r=d['roofline']; print('${name} value=%.3e frac=%.3f'%(d['value'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()})"
}
ARGS="--points 200000000"
run c3_pipe_P4 BAOBZI_B200_PIPE=1
run c3_nopipe_P4 BAOBZI_B200_PIPE=0
run c3_pipe_P2 BAOBZI_B200_PIPE=1 BAOBZI_B200_P=2
ARGS="--points 100000000 --workload C5"
run c5_pipe BAOBZI_B200_PIPE=1
run c5_nopipe BAOBZI_B200_PIPE=0
