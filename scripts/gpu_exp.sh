run() { name=$1; shift
env "$@" python bench.py ${ARGS} --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_$name.json 2> gpurun_out/var_$name.err
python -c "
import json; d=json.load(open('gpurun_out/var_$name.json')); r=d['roofline']; print('$name value=%.3e ms=%.2f'%(d['value'], d['ms_per_step']), 'frac=%.3f'%r['frac'], {k:round(v,2) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/var_$name.err
}
ARGS="--points 400000000"
for c in 4194304 6291456 8388608 12582912 16777216; do run chunk_$c BAOBZI_B200_CHUNK=$c; done
