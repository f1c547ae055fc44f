#!/bin/bash
# A/B of the persistent tile schedule (UMPC_PERSISTENT=1, default) against one tile per warp (=0): parity suite, then
# the device-resident leg three times each
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short -x 2>&1 | tail -8 > gpurun_out/pytest_gpu.log
for pz in 1 0 1 0; do for i in 1 2; do UMPC_PERSISTENT=$pz python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('persistent $pz ms/step %.4f'%d['ms_per_step'],'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'])
except Exception as e: print('FAILED',t[-2:])"; done; done > gpurun_out/persist_ab.log 2>&1
