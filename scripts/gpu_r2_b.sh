# round 2, GPU call B: new kernels -- parity first, then A/B sweeps (all device-resident, 2e8 points unless noted)
set -o pipefail
B="python bench.py --no-cpu --no-e2e --no-extra --no-parity --steps 3"
run() { # label, env..., -- args
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 $B "$@" > gpurun_out/r2b_$label.json 2> gpurun_out/r2b_$label.err
  python - "$label" <<'PY'
import json,sys
l=sys.argv[1]
try:
    d=json.load(open('gpurun_out/r2b_%s.json'%l)); r=d['roofline']
    print('%-28s %.3e evals/s  %.2f ms/step  eval frac %.3f  stages %s  clk %s' % (l, d['value'], d['ms_per_step'], r['frac'] if r else -1, {k:round(v,2) for k,v in (r or {}).get('stages_ms_per_step',{}).items()}, d['clocks'].get('sm_mhz')))
except Exception as e:
    print(l, 'FAILED', e, open('gpurun_out/r2b_%s.err'%l).read()[-300:])
PY
}
P3="--workload C3 --points 200000000"
run c3_piggy            X=1 -- $P3
run c3_nopiggy          BAOBZI_B200_PIGGY=0 -- $P3
run c3_piggy_chunk16m   X=1 -- $P3 --chunk 16777216
run c3_piggy_chunk4m    X=1 -- $P3 --chunk 4194304
run c3_tzfly_4_8        BAOBZI_B200_K4=4,8,1,1,1 -- $P3
run c3_tzfly_3_12       BAOBZI_B200_K4=3,12,1,1,1 -- $P3
run c3_tzfly_4_12       BAOBZI_B200_K4=4,12,1,0,1 -- $P3
run c3_p3_12            BAOBZI_B200_K4=3,12,0,0,1 -- $P3
run c3_mma_12           BAOBZI_B200_K4=4,12,1,1,0 -- $P3
run c3_l2fetch32        BAOBZI_B200_L2_FETCH=32 -- $P3
P5="--workload C5 --points 100000000"
run c5_o12_piggy        X=1 -- $P5
run c5_o12_nopiggy      BAOBZI_B200_PIGGY=0 -- $P5
run c5_o12_sh_3_8       BAOBZI_B200_K4=3,8,0,0,2 -- $P5
run c5_o12_sh_2_12      BAOBZI_B200_K4=2,12,0,1,2 -- $P5
run c5_o12_sh_3_12      BAOBZI_B200_K4=3,12,1,0,2 -- $P5
run c5_o14_default      X=1 -- $P5 --order 14
run c5_o14_sh_2_8       BAOBZI_B200_K4=2,8,0,1,2 -- $P5 --order 14
run c5_o14_sh_3_8       BAOBZI_B200_K4=3,8,1,0,2 -- $P5 --order 14
run c5_o14_sh_2_12      BAOBZI_B200_K4=2,12,1,0,2 -- $P5 --order 14
run c5_o16_default      X=1 -- $P5 --order 16
run c5_o16_sh_2_8k      BAOBZI_B200_K4=2,8,0,1,2 -- $P5 --order 16
run c5_o16_sh_2_8       BAOBZI_B200_K4=2,8,0,0,2 -- $P5 --order 16
run c5_o16_sh_2_12      BAOBZI_B200_K4=2,12,1,0,2 -- $P5 --order 16
run c2_sorted           X=1 -- --workload C2
run c2_table            BAOBZI_B200_SORTED1D=0 -- --workload C2
cuobjdump -sass baobzi_b200/lib/libbaobzi.so | grep -c UBLKCP > gpurun_out/r2b_ublkcp_count.txt; cat gpurun_out/r2b_ublkcp_count.txt
P4="--workload C4 --points 300000000"
run c4_default          X=1 -- $P4
run c4_l2fetch32        BAOBZI_B200_L2_FETCH=32 -- $P4
run c1_1e8_default      X=1 -- --workload C1 --points 100000000
run c1_1e8_l2fetch32    BAOBZI_B200_L2_FETCH=32 -- --workload C1 --points 100000000
timeout 600 python scripts/host_path_ab.py > gpurun_out/r2a_host_ab.txt 2>&1; cat gpurun_out/r2a_host_ab.txt
