#!/bin/bash
# how does one launch scale with the number of resident warps per SM? (latency- vs throughput-bound once-per-solve code)
mkdir -p gpurun_out
for n in 4736 9472 18944 37888 56832 75776; do
python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e --envs $n 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('envs',$n,'ms/step %.4f'%d['ms_per_step'],'value %.4g'%d['value'])
except Exception as e: print('envs',$n,'FAILED',t[-2:])"
done > gpurun_out/occ_sweep.log 2>&1
