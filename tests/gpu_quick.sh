#!/bin/bash
# quick GPU iteration: parity tests + tpb sweep of the device-resident leg
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -x 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
for t in ${TPBS:-32 64 96 128 160}; do python bench.py --steps 10 --warmup 3 --tpb $t --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
txt=sys.stdin.read()
try:
    d=json.loads(txt.strip().splitlines()[-1]); print('tpb',$t,'value %.4g'%d['value'],'ms/step %.3f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'])
except Exception as e: print('tpb',$t,'FAILED',txt[-400:])"; done > gpurun_out/tpb_sweep.log 2>&1
