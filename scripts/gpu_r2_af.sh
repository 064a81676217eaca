# round 2, GPU call AF: L2 evict-first hint on K4's point gather and slot indices (C3), short and power-capped regimes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_variants_gpu.py -m gpu -x -q -k "variant_bits" 2>&1 | tail -2
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 400 python bench.py --no-cpu --no-e2e --no-extra --no-parity "$@" > gpurun_out/r2af_$label.json 2> gpurun_out/r2af_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2af_%s.json'%l)); r=d['roofline']
    print('%-22s %.4e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s %sW' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons'), d['clocks'].get('power_w_max')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2af_%s.err'%l).read()[-300:])
PY
}
run base_2e8      X=1 -- --points 200000000 --steps 10
run evict_2e8     BAOBZI_B200_K4_EVICT=1 -- --points 200000000 --steps 10
run base_1e9      X=1 -- --steps 12
run evict_1e9     BAOBZI_B200_K4_EVICT=1 -- --steps 12
run base_1e9_b    X=1 -- --steps 12
run evict_1e9_b   BAOBZI_B200_K4_EVICT=1 -- --steps 12
export BAOBZI_B200_K4_EVICT=1
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2af_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s 20 -c 30 --csv --log-file gpurun_out/r2af_launches_c3_evict.csv $CMD > gpurun_out/r2af_ncu1.log 2>&1
unset BAOBZI_B200_K4_EVICT
$CMD > gpurun_out/r2af_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s 20 -c 30 --csv --log-file gpurun_out/r2af_launches_c3_base.csv $CMD > gpurun_out/r2af_ncu2.log 2>&1
python - <<'PY'
import csv,collections
for tag in ('base','evict'):
    rows=[r for r in csv.reader(open('gpurun_out/r2af_launches_c3_%s.csv'%tag)) if len(r)>10]
    head=rows[0]; ci={h:i for i,h in enumerate(head)}
    agg=collections.defaultdict(lambda: collections.defaultdict(list))
    for r in rows[1:]:
        name=r[ci['Kernel Name']][:40]; agg[name][r[ci['Metric Name']]].append(float(r[ci['Metric Value']].replace(',','')))
    for n,m in agg.items():
        print(tag, n, {k:(len(v), round(sum(v)/len(v),1)) for k,v in m.items()})
PY
