# round 2, closing GPU call: the full GPU suite, smoke and both bench arms on the last committed build
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2fin3_suite.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -1
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2fin3_bench_ref.json 2> gpurun_out/r2fin3_bench_ref.err; echo "ref rc $?"
timeout 900 python bench.py > gpurun_out/r2fin3_bench.json 2> gpurun_out/r2fin3_bench.err; echo "bench rc $?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2fin3_bench.json'))
print('C3', d['value'], d['ms_per_step'], d['roofline']['frac'], 'e2e', d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks']['reasons'], d.get('gpu_launches'))
for k,v in d.get('workloads',{}).items(): print(k, v.get('value'), (v.get('roofline') or {}).get('frac'), (v.get('parity') or {}).get('max_ulp'))
PY
