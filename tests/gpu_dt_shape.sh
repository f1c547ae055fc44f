#!/bin/bash
# launch-shape sweep of dt_rollout_kernel: resident threads per SM capped through a dynamic shared-memory reservation
mkdir -p gpurun_out; : > gpurun_out/dt_shape.log
for N in ${NS:-10 3 20}; do
for cfg in "128 0" "128 113" "64 113" "64 75" "64 56" "32 113" "32 75" "32 56" "32 45" "32 37"; do
  set -- $cfg
  echo -n "N=$N tpb=$1 smem_kb=$2 : " >> gpurun_out/dt_shape.log
  DT_TPB=$1 DT_SMEM_KB=$2 timeout 300 python bench.py --workload dirtran --horizon $N --steps 3 --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('ms/launch %.3f'%d['ms_per_step'],'env-steps/s %.4g'%d['value'],'iters/solve %.2f'%d['iters_per_solve'])
except Exception as e: print('FAILED',t[-2:])" >> gpurun_out/dt_shape.log
done; done
