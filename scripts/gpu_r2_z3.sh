# round 2, GPU call Z3: 3D order-6 trees of 30 K - 300 K leaves (just above L2) through the binned path and the bucketed path
mkdir -p gpurun_out
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 200 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 2 "$@" > gpurun_out/r2z3_$label.json 2> gpurun_out/r2z3_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2z3_%s.json'%l)); r=d['roofline']; t=d['config']['tree']
    print('%-22s %.3e evals/s  %.3f ms/step  leaves %d  replica %.0f MB  stages %s' % (l, d['value'], d['ms_per_step'], t['leaves'], t['device_bytes']/1e6, {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2z3_%s.err'%l).read()[-300:])
PY
}
for tol in 1e-11 1e-12; do
  T="--workload C3 --order 6 --points 200000000 --tol $tol"
  run t${tol}_binned    X=1 -- $T --mode binned
  run t${tol}_bucketed  X=1 -- $T --mode bucketed
  run t${tol}_gather    BAOBZI_B200_BUCKET_STAGE=0 -- $T --mode bucketed
done
