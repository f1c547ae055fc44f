#!/bin/bash
# round 2, call b: persistent rollout kernel + hybrid TMEM/shared scratch variants (build_variants/), steps per launch, stagger
mkdir -p gpurun_out
O=gpurun_out/r02b_variants.log; : > $O
run() { # label, env..., -- bench args
  local label=$1; shift
  env "$@" 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('$label','ms/step %.4f'%d['ms_per_step'],'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'])
except Exception as e: print('$label','FAILED',t[-1][-200:])" >> $O
}
B="python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rollout.py -m gpu -q --tb=short -x 2>&1 | tail -15 > gpurun_out/r02b_pytest_default.log
UMPC_B200_LIB=$PWD/build_variants/libumpc_b200_m7.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rollout.py -m gpu -q --tb=short -x 2>&1 | tail -15 > gpurun_out/r02b_pytest_m7.log
run "m0 spl1" $B
run "m0 spl1 legacy128" $B --tpb 128
run "m0 spl1 info" UMPC_INFO_EVERY_STEP=1 $B
run "m0 spl64" $B --steps-per-launch 64
run "m0 spl64 75776" $B --steps-per-launch 64 --envs 75776
for m in m4 m6 m7 m8e m9; do
  L=$PWD/build_variants/libumpc_b200_$m.so
  run "$m spl1" UMPC_B200_LIB=$L $B
  run "$m spl64" UMPC_B200_LIB=$L $B --steps-per-launch 64
done
for st in 20 40 60; do
  run "m0 spl64 stagger$st" UMPC_STAGGER_US=$st $B --steps-per-launch 64
  run "m7 spl64 stagger$st" UMPC_B200_LIB=$PWD/build_variants/libumpc_b200_m7.so UMPC_STAGGER_US=$st $B --steps-per-launch 64
done
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_bw tests/microbench/tmem_bw.cu && /tmp/tmem_bw > gpurun_out/r02b_tmem_bw.log 2>&1
