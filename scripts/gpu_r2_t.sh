# round 2, GPU call T: evict-first hint on the tree stream; the new defaults (8192 buckets, 4 Mi-point chunks) without any knob
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_scale_gpu.py -m gpu -x -q -k "bucket or config4" 2>&1 | tail -4
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2t_$label.json 2> gpurun_out/r2t_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2t_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,3) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2t_%s.err'%l).read()[-300:])
PY
}
C4="--workload C4 --points 300000000"
run default        X=1 -- $C4
run default_1e9    X=1 -- --workload C4
run split1         BAOBZI_B200_SORT_SPLIT=1 -- $C4
run c_3m           X=1 -- $C4 --chunk 3145728
run c_5m           X=1 -- $C4 --chunk 5242880
run c_8m           X=1 -- $C4 --chunk 8388608
run p2_4m          BAOBZI_B200_BUCKET_PIPE=2 -- $C4
CMD="python bench.py --workload C4 --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2t_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s 40 -c 60 --csv --log-file gpurun_out/r2t_launches_c4.csv $CMD > gpurun_out/r2t_ncu1.log 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/r2t_launches_c4.csv')) if len(r)>10]
head=rows[0]; ci={h:i for i,h in enumerate(head)}
agg=collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[1:]:
    name=r[ci['Kernel Name']][:60]; agg[name][r[ci['Metric Name']]].append(float(r[ci['Metric Value']].replace(',','')))
for n,m in agg.items():
    print(n, {k:(len(v), round(sum(v)/len(v),1)) for k,v in m.items()})
PY
