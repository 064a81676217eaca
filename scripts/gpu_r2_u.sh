# round 2, GPU call U: chunk size x stages of the pipelined bucket kernel (C4, 3e8 points)
mkdir -p gpurun_out
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2u_$label.json 2> gpurun_out/r2u_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2u_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2u_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
for c in 4718592 5242880 5767168 6291456 7340032; do
  run p3_$c X=1 -- $C4 --chunk $c
  run p2_$c BAOBZI_B200_BUCKET_PIPE=2 -- $C4 --chunk $c
done
run p2_6m_b4k  BAOBZI_B200_BUCKET_PIPE=2 BAOBZI_B200_BUCKETS=4096 -- $C4 --chunk 6291456
