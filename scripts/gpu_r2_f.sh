# round 2, GPU call F: final candidate -- whole GPU suite, smoke, thread split of the sort launches, both bench arms
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r2f_pytest.log 2>&1; tail -12 gpurun_out/r2f_pytest.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/r2f_smoke.log 2>&1; tail -1 gpurun_out/r2f_smoke.log
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2f_$label.json 2> gpurun_out/r2f_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2f_%s.json'%l)); r=d['roofline']
    print('%-28s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'] if r else -1, {k:round(v,3) for k,v in (r or {}).get('stages_ms_per_step',{}).items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2f_%s.err'%l).read()[-300:])
PY
}
P3="--workload C3 --points 200000000"
run c3_default_640_384  X=1 -- $P3
run c3_704_320          BAOBZI_B200_SPLIT_ATHREADS=704 BAOBZI_B200_SPLIT_BTHREADS=320 -- $P3
run c3_768_384          BAOBZI_B200_SPLIT_ATHREADS=768 BAOBZI_B200_SPLIT_BTHREADS=384 -- $P3
run c3_640_448          BAOBZI_B200_SPLIT_ATHREADS=640 BAOBZI_B200_SPLIT_BTHREADS=448 -- $P3
run c3_576_448          BAOBZI_B200_SPLIT_ATHREADS=576 BAOBZI_B200_SPLIT_BTHREADS=448 -- $P3
run c3_640_384_chunk6m  X=1 -- $P3 --chunk 6291456
run c5_o14              X=1 -- --workload C5 --points 100000000 --order 14
run c5_o16              X=1 -- --workload C5 --points 100000000 --order 16
timeout 1200 python bench.py --steps 20 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; cut -c1-200 gpurun_out/r2f_bench.json; tail -3 gpurun_out/r2f_bench.err
timeout 1200 python bench.py > gpurun_out/r2f_bench_steps5.json 2> gpurun_out/r2f_bench_steps5.err; cut -c1-200 gpurun_out/r2f_bench_steps5.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2f_bench_ref.json 2> gpurun_out/r2f_bench_ref.err; cut -c1-200 gpurun_out/r2f_bench_ref.json
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2f_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2f_launches_c3_32Mpts.csv $CMD > gpurun_out/r2f_ncu1.log 2>&1
tail -1 gpurun_out/r2f_ncu1.log
