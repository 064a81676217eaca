# round 2, GPU call O: staged bucket kernel with / without the CTA-local leaf sort, more buckets; two-leaf 1D variants;
# ncu of the staged kernels
mkdir -p gpurun_out
BAOBZI_B200_BUCKET_SORT=0 timeout 900 python -m pytest tests/test_variants_gpu.py -m gpu -x -q -k "bucket" 2>&1 | tail -4
timeout 600 python -m pytest tests/test_variants_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "two_leaf or 1d or golden" 2>&1 | tail -4
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2o_$label.json 2> gpurun_out/r2o_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2o_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2o_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
S="BAOBZI_B200_BUCKET_STAGE=2"
run s_8m_b8k       $S BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 8388608
run s_16m_b8k      $S BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 16777216
run s_8m_b16k      $S BAOBZI_B200_BUCKETS=16384 -- $C4 --chunk 8388608
run s_16m_b16k     $S BAOBZI_B200_BUCKETS=16384 -- $C4 --chunk 16777216
run s_4m_b16k      $S BAOBZI_B200_BUCKETS=16384 -- $C4 --chunk 4194304
run d_8m_b4k       $S BAOBZI_B200_BUCKET_SORT=0 -- $C4 --chunk 8388608
run d_4m_b8k       $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 4194304
run d_8m_b8k       $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 8388608
run d_16m_b8k      $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=8192 -- $C4 --chunk 16777216
run d_4m_b16k      $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=16384 -- $C4 --chunk 4194304
run d_8m_b16k      $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=16384 -- $C4 --chunk 8388608
run d_16m_b16k     $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=16384 -- $C4 --chunk 16777216
run d_8m_b32k      $S BAOBZI_B200_BUCKET_SORT=0 BAOBZI_B200_BUCKETS=32768 -- $C4 --chunk 8388608
for v in 40 41 50 51 31; do
  run c2_v$v BAOBZI_B200_TWOLEAF1D_VARIANT=$v -- --workload C2
done
# ---- ncu: the two staged kernels, one launch each (8 Mi-point chunks, 8192 buckets)
CMD="python bench.py --workload C4 --points 33554432 --chunk 8388608 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
export BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=8192
$CMD > gpurun_out/r2o_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:eval_bucket_staged_kernel -s 4 -c 1 -f -o gpurun_out/r2o_prof_stage_sorted $CMD > gpurun_out/r2o_ncu1.log 2>&1
tail -1 gpurun_out/r2o_ncu1.log
export BAOBZI_B200_BUCKET_SORT=0
$CMD > gpurun_out/r2o_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:eval_bucket_staged_direct -s 4 -c 1 -f -o gpurun_out/r2o_prof_stage_direct $CMD > gpurun_out/r2o_ncu2.log 2>&1
tail -1 gpurun_out/r2o_ncu2.log; ls -la gpurun_out/*.ncu-rep
