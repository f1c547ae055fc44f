#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/cascade.log
for w in cascade-wlcon cascade-wlqp; do
for n in 65536 262144; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --workload $w --envs $n 2>&1 | python -c "
import sys,json
txt=sys.stdin.read()
try:
    d=json.loads(txt.strip().splitlines()[-1]); print('$w',$n,'value %.4g'%d['value'],'ms %.3f'%d['ms_per_step'],d['tracking'])
except Exception: print('$w',$n,'FAILED',txt[-400:])" >> gpurun_out/cascade.log
done; done
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --envs 262144 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('mc 262144 value %.4g'%d['value'],'ms %.3f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'])" >> gpurun_out/cascade.log
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --envs 1048576 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('mc 1048576 value %.4g'%d['value'],'ms %.3f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'])" >> gpurun_out/cascade.log
