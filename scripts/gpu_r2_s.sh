# round 2, GPU call S: launch list (per-kernel time and DRAM bytes, serialised) of the C4 path at 4 Mi-point chunks, 8192 buckets
mkdir -p gpurun_out
export BAOBZI_B200_BUCKET_STAGE=2 BAOBZI_B200_BUCKETS=8192
CMD="python bench.py --workload C4 --points 33554432 --chunk 4194304 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2s_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s 40 -c 60 --csv --log-file gpurun_out/r2s_launches_c4_4m_b8k.csv $CMD > gpurun_out/r2s_ncu1.log 2>&1
tail -2 gpurun_out/r2s_ncu1.log
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/r2s_launches_c4_4m_b8k.csv')) if len(r)>10]
head=rows[0]; ci={h:i for i,h in enumerate(head)}
agg=collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[1:]:
    name=r[ci['Kernel Name']][:60]; agg[name][r[ci['Metric Name']]].append(float(r[ci['Metric Value']].replace(',','')))
for n,m in agg.items():
    print(n, {k:(len(v), round(sum(v)/len(v),1)) for k,v in m.items()})
PY
