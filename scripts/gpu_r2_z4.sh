# round 2, GPU call Z4: 3D order-6 trees between 40 K and 250 K leaves (four-kernel-sort regime of the binned path) against the bucketed path
mkdir -p gpurun_out
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 200 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 2 "$@" > gpurun_out/r2z4_$label.json 2> gpurun_out/r2z4_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2z4_%s.json'%l)); r=d['roofline']; t=d['config']['tree']
    print('%-22s %.3e evals/s  %.3f ms/step  leaves %d  cells %d  replica %.0f MB  stages %s' % (l, d['value'], d['ms_per_step'], t['leaves'], t['grid_cells'], t['device_bytes']/1e6, {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2z4_%s.err'%l).read()[-300:])
PY
}
for tol in 5e-12 2.5e-12; do
  T="--workload C3 --order 6 --points 200000000 --tol $tol"
  run t${tol}_binned    X=1 -- $T --mode binned
  run t${tol}_bucketed  X=1 -- $T --mode bucketed
done
