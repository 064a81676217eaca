# round 2, GPU call C: packed-key split sort, piggy A/B + ncu of both, reworked few-leaf 1D kernel, new high-order defaults,
# host-path A/B, ncu captures for profiles/
timeout 900 python -m pytest tests/test_variants_gpu.py tests/test_scale_gpu.py tests/test_reference_consumers_gpu.py tests/test_parity_gpu.py -m gpu -q --durations=5 -k "not timing_c" > gpurun_out/r2c_pytest.log 2>&1; tail -6 gpurun_out/r2c_pytest.log
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() {
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2c_$label.json 2> gpurun_out/r2c_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2c_%s.json'%l)); r=d['roofline']
    print('%-28s %.3e evals/s  %.3f ms/step  frac %.3f  stages %s  clk %s %s' % (l, d['value'], d['ms_per_step'], r['frac'] if r else -1, {k:round(v,3) for k,v in (r or {}).get('stages_ms_per_step',{}).items()}, d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2c_%s.err'%l).read()[-300:])
PY
}
P3="--workload C3 --points 200000000"
run c3_split_packed     X=1 -- $P3
run c3_piggy_packed     BAOBZI_B200_PIGGY=1 -- $P3
run c3_single_launch    BAOBZI_B200_SORT_SPLIT=0 -- $P3
run c3_1e9              X=1 -- --workload C3
run c2_sorted6          X=1 -- --workload C2
run c2_sorted5          BAOBZI_B200_SORTED1D_CTAS=5 -- --workload C2
run c2_table            BAOBZI_B200_SORTED1D=0 -- --workload C2
P5="--workload C5 --points 100000000"
run c5_o12              X=1 -- $P5
run c5_o14_shared       X=1 -- $P5 --order 14
run c5_o14_runwarp      BAOBZI_B200_K4=2,8,0,1,1 -- $P5 --order 14
run c5_o16_shared       X=1 -- $P5 --order 16
run c5_o16_runwarp      BAOBZI_B200_K4=2,4,0,1,1 -- $P5 --order 16
run c4                  X=1 -- --workload C4 --points 300000000
run c1_1e8              X=1 -- --workload C1 --points 100000000
timeout 900 python scripts/host_path_ab.py > gpurun_out/r2c_host_ab.txt 2>&1; cat gpurun_out/r2c_host_ab.txt
# ---- ncu (one tool per call): launch list, then full captures of the kernels the rooflines quote
CMD="python bench.py --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD > gpurun_out/r2c_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2c_launches_c3_32Mpts.csv $CMD > gpurun_out/r2c_ncu1.log 2>&1
$CMD > gpurun_out/r2c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_eval3d_kernel -s 6 -c 1 -f -o gpurun_out/r2c_prof_k4 $CMD > gpurun_out/r2c_ncu2.log 2>&1
$CMD > gpurun_out/r2c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bin_sort_fused -s 6 -c 2 -f -o gpurun_out/r2c_prof_sort $CMD > gpurun_out/r2c_ncu3.log 2>&1
export BAOBZI_B200_PIGGY=1
$CMD > gpurun_out/r2c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:piggy -s 2 -c 1 -f -o gpurun_out/r2c_prof_piggy $CMD > gpurun_out/r2c_ncu4.log 2>&1
unset BAOBZI_B200_PIGGY
CMD2="python bench.py --workload C2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD2 > gpurun_out/r2c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:eval_sorted_1d -s 3 -c 1 -f -o gpurun_out/r2c_prof_sorted1d $CMD2 > gpurun_out/r2c_ncu5.log 2>&1
CMD3="python bench.py --workload C5 --order 16 --points 33554432 --steps 1 --warmup 3 --no-cpu --no-e2e --no-profile --no-extra --no-parity"
$CMD3 > gpurun_out/r2c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:shared_kernel -s 3 -c 1 -f -o gpurun_out/r2c_prof_shared16 $CMD3 > gpurun_out/r2c_ncu6.log 2>&1
tail -2 gpurun_out/r2c_ncu6.log; ls -la gpurun_out/*.ncu-rep
