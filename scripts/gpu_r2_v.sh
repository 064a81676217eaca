# round 2, GPU call V: both sort launches beside the evaluation (bucketed path); new defaults (5.5 Mi-point chunks, two stages)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_scale_gpu.py -m gpu -x -q -k "bucket or config4" 2>&1 | tail -4
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2v_$label.json 2> gpurun_out/r2v_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2v_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2v_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
run default        X=1 -- $C4
run overlap        BAOBZI_B200_BUCKET_OVERLAP=1 -- $C4
run overlap_5m     BAOBZI_B200_BUCKET_OVERLAP=1 -- $C4 --chunk 5242880
run overlap_4m     BAOBZI_B200_BUCKET_OVERLAP=1 -- $C4 --chunk 4194304
run overlap_6m     BAOBZI_B200_BUCKET_OVERLAP=1 -- $C4 --chunk 6291456
run overlap_1e9    BAOBZI_B200_BUCKET_OVERLAP=1 -- --workload C4
