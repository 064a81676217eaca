# round 2, GPU call AC: 2D bucket kernels evaluate a thread's two points of a batch side by side
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_scale_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "bucket or config4 or random" 2>&1 | tail -3
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2ac_$label.json 2> gpurun_out/r2ac_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2ac_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.4f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,4) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2ac_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
run pair             X=1 -- $C4
run pair_6m          X=1 -- $C4 --chunk 6291456
run pair_7m          X=1 -- $C4 --chunk 7340032
run pair_p3          BAOBZI_B200_BUCKET_PIPE=3 -- $C4
