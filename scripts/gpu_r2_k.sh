# round 2, GPU call K: lookup-launch points in flight (carve-out fixed), C4 random-access floor
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2k_$label.json 2> gpurun_out/r2k_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2k_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2k_%s.err'%l).read()[-300:])
PY
}
P3="--workload C3 --points 200000000"
run c3_default            X=1 -- $P3
run c3_apts8_448_448      BAOBZI_B200_SPLIT_APTS=8 BAOBZI_B200_SPLIT_ATHREADS=448 BAOBZI_B200_SPLIT_BTHREADS=448 -- $P3
run c3_apts8_512_384      BAOBZI_B200_SPLIT_APTS=8 BAOBZI_B200_SPLIT_ATHREADS=512 BAOBZI_B200_SPLIT_BTHREADS=384 -- $P3
run c3_apts2_768_448      BAOBZI_B200_SPLIT_APTS=2 BAOBZI_B200_SPLIT_ATHREADS=768 BAOBZI_B200_SPLIT_BTHREADS=448 -- $P3
run c3_apts2_640_448      BAOBZI_B200_SPLIT_APTS=2 BAOBZI_B200_SPLIT_ATHREADS=640 BAOBZI_B200_SPLIT_BTHREADS=448 -- $P3
scripts/micro/_bin/c4_floor > gpurun_out/r2k_c4_floor.txt 2>&1; cat gpurun_out/r2k_c4_floor.txt
