# round 2, GPU call L: C4 chunk size (the floor micro-benchmark says L2-resident chunks cut the random accesses 4x)
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 --workload C4 --points 300000000"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2l_$label.json 2> gpurun_out/r2l_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2l_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2l_%s.err'%l).read()[-300:])
PY
}
run c4_chunk16m   X=1 --
run c4_chunk8m    X=1 -- --chunk 8388608
run c4_chunk4m    X=1 -- --chunk 4194304
run c4_chunk2m    X=1 -- --chunk 2097152
run c4_chunk4m_b16k  BAOBZI_B200_BUCKETS=16384 -- --chunk 4194304
run c4_chunk4m_b1k   BAOBZI_B200_BUCKETS=1024 -- --chunk 4194304
