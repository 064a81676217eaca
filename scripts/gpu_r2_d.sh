# round 2, GPU call D: two-leaf 1D kernel, piggyback side jobs rewritten for ILP, then the full default bench
timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_scale_gpu.py -m gpu -q --durations=5 > gpurun_out/r2d_pytest.log 2>&1; tail -6 gpurun_out/r2d_pytest.log
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2d_$label.json 2> gpurun_out/r2d_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2d_%s.json'%l)); r=d['roofline']
    print('%-28s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'] if r else -1, {k:round(v,3) for k,v in (r or {}).get('stages_ms_per_step',{}).items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2d_%s.err'%l).read()[-300:])
PY
}
P3="--workload C3 --points 200000000"
run c2_twoleaf          X=1 -- --workload C2
run c2_sorted5          BAOBZI_B200_TWOLEAF1D=0 BAOBZI_B200_SORTED1D_CTAS=5 -- --workload C2
run c2_table            BAOBZI_B200_TWOLEAF1D=0 BAOBZI_B200_SORTED1D=0 -- --workload C2
run c3_split            X=1 -- $P3
run c3_piggy_ilp        BAOBZI_B200_PIGGY=1 -- $P3
run c3_piggy_ilp_4m     BAOBZI_B200_PIGGY=1 -- $P3 --chunk 4194304
run c3_split_16m        X=1 -- $P3 --chunk 16777216
run c5_piggy_ilp        BAOBZI_B200_PIGGY=1 -- --workload C5 --points 100000000
timeout 1200 python bench.py --steps 20 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; cut -c1-300 gpurun_out/r2d_bench.json; tail -3 gpurun_out/r2d_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2d_bench_ref.json 2> gpurun_out/r2d_bench_ref.err; cut -c1-200 gpurun_out/r2d_bench_ref.json
export BAOBZI_B200_PIGGY=1
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2d_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:piggy -s 2 -c 1 -f -o gpurun_out/r2d_prof_piggy $CMD > gpurun_out/r2d_ncu1.log 2>&1
unset BAOBZI_B200_PIGGY
CMD2="python bench.py --workload C2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD2 > gpurun_out/r2d_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:two_leaf -s 3 -c 1 -f -o gpurun_out/r2d_prof_twoleaf $CMD2 > gpurun_out/r2d_ncu2.log 2>&1
tail -1 gpurun_out/r2d_ncu2.log
