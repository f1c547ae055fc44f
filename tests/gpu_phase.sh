#!/bin/bash
# time of the set-up phase alone / set-up + loop / whole update (all warps start the phase together)
mkdir -p gpurun_out
for ph in 1 2 0; do for n in 18944 65536; do
UMPC_DEBUG_PHASE=$ph python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e --envs $n 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('phase',$ph,'envs',$n,'ms/step %.4f'%d['ms_per_step'])
except Exception as e: print('phase',$ph,'envs',$n,'FAILED',t[-2:])"
done; done > gpurun_out/phase_sweep.log 2>&1
