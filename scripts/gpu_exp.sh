cp baobzi_b200/lib/libbaobzi.so /tmp/good.so
for e in 3 4; do
  cp baobzi_b200/lib/libbaobzi_exp$e.so baobzi_b200/lib/libbaobzi.so
  python bench.py --points 200000000 --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_sortexp$e.json 2> gpurun_out/var_sortexp$e.err
  python -c "
import json; d=json.load(open('gpurun_out/var_sortexp$e.json')); print('sortexp$e', {k:round(v,2) for k,v in d['roofline']['stages_ms_per_step'].items()})" || tail -3 gpurun_out/var_sortexp$e.err
done
cp /tmp/good.so baobzi_b200/lib/libbaobzi.so
