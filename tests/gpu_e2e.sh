#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/e2e_sweep.log
for n in 1 2 3 4 6; do
UMPC_E2E_CHUNKS=$n timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunks',$n,'e2e %.4g'%d['e2e']['value'],'ms %.3f'%d['e2e']['ms_per_step'],'value %.4g'%d['value'])" >> gpurun_out/e2e_sweep.log
done
