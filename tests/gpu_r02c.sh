#!/bin/bash
# round 2, call c: where did the single-step launch lose 30%?  steps per launch x info x old/new library
mkdir -p gpurun_out
O=gpurun_out/r02c_ab.log; : > $O
run() { local label=$1; shift
  env "$@" 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('$label','ms/envstep %.4f'%(d['ms_per_step']/d['plant_steps_per_launch']),'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'])
except Exception as e: print('$label','FAILED',t[-1][-200:])" >> $O
}
B="python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e"
OLD=$PWD/build_variants/libumpc_b200_old.so
for spl in 1 2 4 16 64; do
  run "new spl$spl" $B --steps-per-launch $spl
  run "new spl$spl info" UMPC_INFO_EVERY_STEP=1 $B --steps-per-launch $spl
  run "new spl$spl tpb128" $B --steps-per-launch $spl --tpb 128
  run "old spl$spl" UMPC_B200_LIB=$OLD $B --steps-per-launch $spl
done
