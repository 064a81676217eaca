# GPU-box experiment: K4 variants (points per lane P, smem staging) x chunk size on workload C3
python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "ragged or device_pointers" > gpurun_out/pytest_var.log 2>&1; tail -2 gpurun_out/pytest_var.log
for P in 2 4; do for S in 0 1; do for CH in 2097152 8388608; do
  BAOBZI_B200_P=$P BAOBZI_B200_SMEM=$S BAOBZI_B200_CHUNK=$CH python bench.py --points 200000000 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_${P}_${S}_${CH}.json 2> gpurun_out/var.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/var_${P}_${S}_${CH}.json')); r=d['roofline']; print('P=$P smem=$S chunk=$CH value=%.3e frac=%.3f'%(d['value'], r['frac']), {k:round(v,2) for k,v in r['stages_ms_per_step'].items()})"
done; done; done
