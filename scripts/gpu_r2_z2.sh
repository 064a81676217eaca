# round 2, GPU call Z2: the 1 M-leaf 3D tree (1.8 GB) through the binned path (auto) and through the bucketed path
mkdir -p gpurun_out
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 240 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 2 "$@" > gpurun_out/r2z2_$label.json 2> gpurun_out/r2z2_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2z2_%s.json'%l)); r=d['roofline']; t=d['config']['tree']
    print('%-22s %.3e evals/s  %.3f ms/step  leaves %d  replica %.0f MB  stages %s' % (l, d['value'], d['ms_per_step'], t['leaves'], t['device_bytes']/1e6, {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2z2_%s.err'%l).read()[-300:])
PY
}
T="--workload C3 --order 6 --points 100000000 --tol 1e-13"
run bucketed          X=1 -- $T --mode bucketed
run bucketed_4m       X=1 -- $T --mode bucketed --chunk 4194304
run binned            X=1 -- $T --mode binned
