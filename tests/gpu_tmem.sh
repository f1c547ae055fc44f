#!/bin/bash
mkdir -p gpurun_out
for t in ${TEST_TPBS:-256}; do
  UMPC_TPB=$t timeout 300 python -m pytest tests -m gpu -q --tb=line -x 2>&1 | tail -4 > gpurun_out/pytest_gpu_tpb$t.log
done
: > gpurun_out/variants.log
for lib in "" build_variants/*.so; do
for t in ${TPBS:-256 128}; do
  [ -n "$lib" ] && [ ! -f "$lib" ] && continue
  UMPC_B200_LIB=${lib:+$PWD/$lib} timeout 120 python bench.py --steps 10 --warmup 3 --tpb $t --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
txt=sys.stdin.read()
try:
    d=json.loads(txt.strip().splitlines()[-1]); print('${lib:-default}','tpb',$t,'value %.4g'%d['value'],'ms/step %.3f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'],'nonfinite',d['tracking']['non_finite_envs'],'err %.4f'%d['tracking']['mean_err_mm2'])
except Exception as e: print('${lib:-default}','tpb',$t,'FAILED',txt[-300:])" >> gpurun_out/variants.log
done
done
