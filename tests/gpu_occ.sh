#!/bin/bash
mkdir -p gpurun_out
for pad in 0 80 150 270 600; do UMPC_SMEM_PAD_FLOATS=$pad python bench.py --steps 10 --warmup 3 --tpb 32 --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pad',$pad,'value %.4g'%d['value'],'ms/step %.3f'%d['ms_per_step'])"; done > gpurun_out/occ.log 2>&1
