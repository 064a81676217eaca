# round 2, GPU call E: coalescing scatter of the split sort (rank in the key -> CTA-local order for free)
timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_scale_gpu.py tests/test_parity_gpu.py -m gpu -q --durations=3 > gpurun_out/r2e_pytest.log 2>&1; tail -5 gpurun_out/r2e_pytest.log
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2e_$label.json 2> gpurun_out/r2e_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2e_%s.json'%l)); r=d['roofline']
    print('%-28s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'] if r else -1, {k:round(v,3) for k,v in (r or {}).get('stages_ms_per_step',{}).items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2e_%s.err'%l).read()[-300:])
PY
}
P3="--workload C3 --points 200000000"
run c3_coal             X=1 -- $P3
run c3_nocoal           BAOBZI_B200_SPLIT_COAL=0 -- $P3
run c3_coal_athr768     BAOBZI_B200_SPLIT_ATHREADS=768 BAOBZI_B200_SPLIT_BTHREADS=256 -- $P3
run c3_coal_athr640     BAOBZI_B200_SPLIT_ATHREADS=640 BAOBZI_B200_SPLIT_BTHREADS=384 -- $P3
run c1_1e8_coal         X=1 -- --workload C1 --points 100000000
run c1_1e8_nocoal       BAOBZI_B200_SPLIT_COAL=0 -- --workload C1 --points 100000000
run c5_coal_forced      BAOBZI_B200_SORT_SPLIT=1 -- --workload C5 --points 100000000
run c5_default          X=1 -- --workload C5 --points 100000000
run c3_1e9              X=1 -- --workload C3
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2e_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2e_launches_c3_32Mpts.csv $CMD > gpurun_out/r2e_ncu1.log 2>&1
$CMD > gpurun_out/r2e_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_sort_fused -s 6 -c 2 -f -o gpurun_out/r2e_prof_sort $CMD > gpurun_out/r2e_ncu2.log 2>&1
tail -1 gpurun_out/r2e_ncu2.log
