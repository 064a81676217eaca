# round 2, GPU call X: full suite + smoke + both bench arms on the committed build; C1 through the bucketed / staged path;
# C4 sort knobs; ncu of the final pipelined bucket kernel
mkdir -p gpurun_out
echo "== full gpu suite =="
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2x_suite.txt
echo "== smoke =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2
echo "== bench reference arm =="
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2x_bench_ref.json 2> gpurun_out/r2x_bench_ref.err; tail -c 300 gpurun_out/r2x_bench_ref.json; echo
echo "== bench default =="
timeout 900 python bench.py > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2x_bench.json'))
print('C3', d['value'], d['ms_per_step'], d['roofline']['frac'], 'e2e', d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks']['reasons'])
for k,v in d.get('workloads',{}).items(): print(k, v.get('value'), (v.get('roofline') or {}).get('frac'), (v.get('roofline') or {}).get('traffic_source'), (v.get('parity') or {}).get('max_ulp'))
PY
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3 "$@" > gpurun_out/r2x_$label.json 2> gpurun_out/r2x_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2x_%s.json'%l)); r=d['roofline']
    print('%-26s %.3e evals/s  %.4f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,4) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2x_%s.err'%l).read()[-300:])
PY
}
echo "== C1 (2D order 8, 8860 leaves, 5 MB tree) through the bucketed path =="
run c1_1e6_auto     X=1 -- --workload C1 --steps 200
for b in 148 296 512 1024 4096; do
  run c1_1e6_bkt$b  BAOBZI_B200_BUCKETS=$b -- --workload C1 --steps 200 --mode bucketed
done
run c1_1e6_direct   X=1 -- --workload C1 --steps 200 --mode direct
run c1_1e8_auto     X=1 -- --workload C1 --points 100000000
for b in 512 1024 4096; do
  run c1_1e8_bkt$b  BAOBZI_B200_BUCKETS=$b -- --workload C1 --points 100000000 --mode bucketed
done
echo "== C4 sort knobs =="
C4="--workload C4 --points 300000000"
run c4_default      X=1 -- $C4
run c4_split2       BAOBZI_B200_SORT_SPLIT=2 -- $C4
run c4_a768_b256    BAOBZI_B200_SPLIT_ATHREADS=768 BAOBZI_B200_SPLIT_BTHREADS=256 -- $C4
run c4_a384_b512    BAOBZI_B200_SPLIT_ATHREADS=384 BAOBZI_B200_SPLIT_BTHREADS=512 -- $C4
run c4_a256_b512    BAOBZI_B200_SPLIT_ATHREADS=256 BAOBZI_B200_SPLIT_BTHREADS=512 -- $C4
echo "== ncu: final pipelined bucket kernel =="
CMD="python bench.py --workload C4 --points 34603008 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2x_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:eval_bucket_pipe_kernel -s 8 -c 1 -f -o gpurun_out/r2x_prof_pipe $CMD > gpurun_out/r2x_ncu1.log 2>&1
tail -1 gpurun_out/r2x_ncu1.log
