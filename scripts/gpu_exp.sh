# GPU-box A/B on ONE box: the split sort (default from 2048 leaves / buckets) against the single cooperative launch
run() { # label, env..., bench args
  label=$1; shift
  timeout 300 env "$@" > gpurun_out/x.json 2> gpurun_out/x.err; python -c "
import json; d=json.load(open('gpurun_out/x.json')); r=d['roofline']; print('$label value=%.4e ms=%.3f'%(d['value'], d['ms_per_step']), {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks']['sm_mhz'], d['clocks']['reasons'])" || tail -3 gpurun_out/x.err
}
for sp in 0 1; do
run "C3 2e8 split=$sp" BAOBZI_B200_SORT_SPLIT=$sp python bench.py --workload C3 --points 200000000 --steps 5 --warmup 3 --no-cpu --no-e2e
run "C1 1e8 split=$sp" BAOBZI_B200_SORT_SPLIT=$sp python bench.py --workload C1 --points 100000000 --steps 5 --warmup 3 --no-cpu --no-e2e
run "C4 3e8 split=$sp" BAOBZI_B200_SORT_SPLIT=$sp python bench.py --workload C4 --points 300000000 --steps 3 --warmup 3 --no-cpu --no-e2e
run "C5 2e8 split=$sp" BAOBZI_B200_SORT_SPLIT=$sp python bench.py --workload C5 --points 200000000 --steps 3 --warmup 3 --no-cpu --no-e2e
done
