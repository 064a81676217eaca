# round 2, GPU call W: overlap schedule with the sort launches on the same shared-memory carve-out as the evaluation kernel
mkdir -p gpurun_out
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2w_$label.json 2> gpurun_out/r2w_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2w_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2w_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
run overlap        BAOBZI_B200_BUCKET_OVERLAP=1 -- $C4
run overlap_6m     BAOBZI_B200_BUCKET_OVERLAP=1 -- $C4 --chunk 6291456
