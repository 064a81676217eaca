# round 2, final GPU call: full suite, smoke, both bench arms on the committed build; the 997 K-leaf 3D tree in auto mode
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2fin_suite.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2fin_bench_ref.json 2> gpurun_out/r2fin_bench_ref.err; tail -c 200 gpurun_out/r2fin_bench_ref.json; echo
timeout 900 python bench.py > gpurun_out/r2fin_bench.json 2> gpurun_out/r2fin_bench.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2fin_bench.json'))
print('C3', d['value'], d['ms_per_step'], d['roofline']['frac'], 'e2e', d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks']['reasons'])
for k,v in d.get('workloads',{}).items(): print(k, v.get('value'), (v.get('roofline') or {}).get('frac'), (v.get('roofline') or {}).get('one_block_gather_per_eval'), (v.get('parity') or {}).get('max_ulp'))
PY
timeout 240 python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 2 --workload C3 --order 6 --points 100000000 --tol 1e-13 > gpurun_out/r2fin_3d_1Mleaves_auto.json 2> gpurun_out/r2fin_3d.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2fin_3d_1Mleaves_auto.json')); r=d['roofline']
print('3D 1M leaves auto', d['value'], r['stages_ms_per_step'])
PY
