# round 2, GPU call I: guard-band memory check, all-GPU host test, single-chunk split experiment (C1 at 1 M points)
timeout 1500 python -m pytest tests/test_guard_bands_gpu.py tests/test_scale_gpu.py tests/test_parity_gpu.py -m gpu -q --durations=5 > gpurun_out/r2i_pytest.log 2>&1; tail -8 gpurun_out/r2i_pytest.log
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2i_$label.json 2> gpurun_out/r2i_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2i_%s.json'%l)); r=d['roofline']
    print('%-22s %.3e evals/s  %.4f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'], {k:round(v,4) for k,v in r['stages_ms_per_step'].items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2i_%s.err'%l).read()[-300:])
PY
}
run c1_1m_default     X=1 -- --workload C1 --steps 10000
run c1_1m_splitsingle BAOBZI_B200_SPLIT_SINGLE=1 -- --workload C1 --steps 10000
run c1_4m_default     X=1 -- --workload C1 --points 4000000 --steps 3000
run c1_4m_splitsingle BAOBZI_B200_SPLIT_SINGLE=1 -- --workload C1 --points 4000000 --steps 3000
run c3_8m_default     X=1 -- --workload C3 --points 8000000 --steps 1000
run c3_8m_splitsingle BAOBZI_B200_SPLIT_SINGLE=1 -- --workload C3 --points 8000000 --steps 1000
cat gpurun_out/guard_bands_piggy0.txt | tail -4
