# round 2, GPU call N: bucketed path with the bucket's piece of the tree staged in shared memory by TMA (bucket_stage.cu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_variants_gpu.py -m gpu -x -q -k "bucket" 2>&1 | tail -8
timeout 900 python -m pytest tests/test_scale_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "config4 or bucketed or random or golden_device" 2>&1 | tail -8
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 --workload C4 --points 300000000"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2n_$label.json 2> gpurun_out/r2n_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2n_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2n_%s.err'%l).read()[-300:])
PY
}
run gather_16m      BAOBZI_B200_BUCKET_STAGE=0 -- --chunk 16777216
run gather_4m       BAOBZI_B200_BUCKET_STAGE=0 -- --chunk 4194304
run stage_16m       BAOBZI_B200_BUCKET_STAGE=2 -- --chunk 16777216
run stage_8m        BAOBZI_B200_BUCKET_STAGE=2 -- --chunk 8388608
run stage_4m        BAOBZI_B200_BUCKET_STAGE=2 -- --chunk 4194304
run stage_2m        BAOBZI_B200_BUCKET_STAGE=2 -- --chunk 2097152
run stage_4m_b8k    BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=8192 -- --chunk 4194304
run stage_8m_b8k    BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=8192 -- --chunk 8388608
run stage_4m_b2k    BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=2048 -- --chunk 4194304
run stage_2m_b8k    BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=8192 -- --chunk 2097152
