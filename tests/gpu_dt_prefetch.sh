#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/dt_prefetch.log
timeout 600 python -m pytest tests/test_gpu_dirtran.py -q --tb=short 2>&1 | tail -5 >> gpurun_out/dt_prefetch.log
for N in 3 5 10 15 20; do for pf in 0 1; do
  echo -n "N=$N DT_PREFETCH=$pf : " >> gpurun_out/dt_prefetch.log
  DT_PREFETCH=$pf timeout 300 python bench.py --workload dirtran --horizon $N --steps 4 --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('ms/launch %.3f'%d['ms_per_step'],'env-steps/s %.4g'%d['value'],'iters/solve %.2f'%d['iters_per_solve'])
except Exception as e: print('FAILED',t[-2:])" >> gpurun_out/dt_prefetch.log
done; done
