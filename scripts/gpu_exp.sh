python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
run() { name=$1; shift
env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_$name.json 2> gpurun_out/var_$name.err
python -c "
import json; d=json.load(open('gpurun_out/var_$name.json')); r=d['roofline']; print('$name value=%.3e'%d['value'], r.get('kernel'), r['bound'], 'frac=%.3f'%r['frac'], {k:round(v,2) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/var_$name.err
}
ARGS="--workload C1 --points 100000000"; run c1_auto X=1
ARGS="--workload C1 --points 100000000 --mode binned --chunk 4194304"; run c1_binned_4m X=1
ARGS="--workload C1"; run c1_1m X=1
