# round 2, GPU call P: staged bucket kernel, CTA-local order by grid cell (lookups after the sort, in sorted order)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_variants_gpu.py -m gpu -x -q -k "bucket or two_leaf" 2>&1 | tail -4
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2p_$label.json 2> gpurun_out/r2p_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2p_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2p_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
S="BAOBZI_B200_BUCKET_STAGE=2"
run bin_8m_b8k     $S BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 8388608
run bin_4m_b8k     $S BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 4194304
run bin_16m_b8k    $S BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 16777216
run bin_8m_b4k     $S -- $C4 --chunk 8388608
run leaf_8m_b8k    $S BAOBZI_B200_BUCKET_SORT=1 BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 8388608
run c2_default     X=1 -- --workload C2
CMD="python bench.py --workload C4 --points 33554432 --chunk 8388608 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
export BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=8192
$CMD > gpurun_out/r2p_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:eval_bucket_staged_kernel -s 4 -c 1 -f -o gpurun_out/r2p_prof_stage_bybin $CMD > gpurun_out/r2p_ncu1.log 2>&1
tail -1 gpurun_out/r2p_ncu1.log
