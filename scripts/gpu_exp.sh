# A/B of the sort variants on one box
run() { # label, env..., bench args
  label=$1; shift
  timeout 300 env "$@" > gpurun_out/x.json 2> gpurun_out/x.err; python -c "
import json; d=json.load(open('gpurun_out/x.json')); r=d['roofline']; print('$label value=%.4e ms=%.3f'%(d['value'], d['ms_per_step']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()})" || tail -3 gpurun_out/x.err
}
for w in 0 1 0 1; do
run "C3 2e8 wait=$w" BAOBZI_B200_SPLIT_WAIT=$w python bench.py --workload C3 --points 200000000 --steps 5 --warmup 3 --no-cpu --no-e2e
run "C1 1e8 wait=$w" BAOBZI_B200_SPLIT_WAIT=$w python bench.py --workload C1 --points 100000000 --steps 5 --warmup 3 --no-cpu --no-e2e
done
