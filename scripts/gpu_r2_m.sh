# round 2, GPU call M: the no-allocation entry point alone in a fresh process, then the whole suite (which ends with
# it), smoke, and both bench arms on the build the driver will see
mkdir -p gpurun_out
echo "== test_zz_no_alloc alone =="
timeout 900 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r2m_noalloc_alone.txt
echo "== full gpu suite =="
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r2m_suite.txt
echo "== smoke =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -3
echo "== bench reference arm =="
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2m_bench_ref.json 2> gpurun_out/r2m_bench_ref.err; tail -c 600 gpurun_out/r2m_bench_ref.json
echo "== bench =="
timeout 900 python bench.py > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2m_bench.json'))
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['clocks'], d.get('parity'))
for k,v in d.get('workloads',{}).items(): print(k, v.get('value'), v.get('roofline',{}).get('frac'), v.get('parity'))
PY
