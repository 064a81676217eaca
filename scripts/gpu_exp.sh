timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -1
run() {
timeout 300 python bench.py --workload $1 --points $2 --steps $3 --warmup 5 --no-cpu --no-e2e > gpurun_out/x.json 2> gpurun_out/x.err
python -c "
import json; d=json.load(open('gpurun_out/x.json')); r=d['roofline']; print('$1 n $2 value=%.4e ms=%.4f'%(d['value'], d['ms_per_step']), {k:round(v,4) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/x.err
}
run C1 1000000 20; run C1 100000000 3; run C3 200000000 3; run C4 300000000 2
