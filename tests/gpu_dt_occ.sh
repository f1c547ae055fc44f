#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/dt_occ.log
timeout 600 python -m pytest tests/test_gpu_dirtran.py -q --tb=short 2>&1 | tail -3 >> gpurun_out/dt_occ.log
for N in 3 5 10 20; do
  echo -n "N=$N : " >> gpurun_out/dt_occ.log
  timeout 300 python bench.py --workload dirtran --horizon $N --steps 4 --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('ms/launch %.3f'%d['ms_per_step'],'env-steps/s %.4g'%d['value'],'iters/solve %.2f'%d['iters_per_solve'])
except Exception as e: print('FAILED',t[-2:])" >> gpurun_out/dt_occ.log
done
