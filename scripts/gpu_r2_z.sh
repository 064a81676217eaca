# round 2, GPU call Z: a 3D tree larger than L2 (C3's function at order 6, tighter tolerance) through every path
mkdir -p gpurun_out
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 400 python bench.py --no-cpu --no-e2e --no-extra --steps 3 "$@" > gpurun_out/r2z_$label.json 2> gpurun_out/r2z_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2z_%s.json'%l)); r=d['roofline']; t=d['config']['tree']
    print('%-22s %.3e evals/s  %.3f ms/step  leaves %d  replica %.0f MB  stages %s  parity %s' % (l, d['value'], d['ms_per_step'], t['leaves'], t['device_bytes']/1e6, {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, (d['config'].get('parity') or {}).get('max_ulp')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2z_%s.err'%l).read()[-300:])
PY
}
T="--workload C3 --order 6 --points 200000000"
run tol1e-13_auto     X=1 -- $T --tol 1e-13
run tol3e-14_auto     X=1 -- $T --tol 3e-14
run tol3e-14_gather   BAOBZI_B200_BUCKET_STAGE=0 -- $T --tol 3e-14
run tol3e-14_binned   X=1 -- $T --tol 3e-14 --mode binned
run tol3e-14_direct   X=1 -- $T --tol 3e-14 --mode direct
run tol3e-14_queue    BAOBZI_B200_BUCKET_PIPE=0 -- $T --tol 3e-14
run tol3e-14_8m       X=1 -- $T --tol 3e-14 --chunk 8388608
run tol3e-14_2m       X=1 -- $T --tol 3e-14 --chunk 2097152
