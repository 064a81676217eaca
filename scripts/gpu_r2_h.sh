# power-capped regime (1e9 points, 10 steps): does less DRAM traffic per point buy clock?
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 10 --workload C3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2h_$label.json 2> gpurun_out/r2h_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2h_%s.json'%l)); r=d['roofline']
    print('%-22s %.3e evals/s  %.2f ms/step  frac %.3f  stages %s  clk %s %s %sW' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,2) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons'), d['clocks'].get('power_w_max')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2h_%s.err'%l).read()[-300:])
PY
}
run default_576_448   X=1 --
run chunk4m           X=1 -- --chunk 4194304
run chunk6m           X=1 -- --chunk 6291456
run l2fetch32         BAOBZI_B200_L2_FETCH=32 --
run l2fetch32_chunk4m BAOBZI_B200_L2_FETCH=32 -- --chunk 4194304
run t_640_384         BAOBZI_B200_SPLIT_ATHREADS=640 BAOBZI_B200_SPLIT_BTHREADS=384 --
run nocoal            BAOBZI_B200_SPLIT_COAL=0 --
