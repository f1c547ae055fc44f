#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r02i_chunks_dynamic.log; : > $O
run() { local label=$1; shift
  env "$@" 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('$label','ms/envstep %.4f'%(d['ms_per_step']/d['plant_steps_per_launch']),'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'])
except Exception as e: print('$label','FAILED',t[-1][-200:])" >> $O
}
B="python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e"
for c in 4 8 16; do run "chunks$c spl125" UMPC_ROLLOUT_CHUNKS=$c $B; done
for c in 8; do run "chunks$c spl128" UMPC_ROLLOUT_CHUNKS=$c $B --steps-per-launch 128; run "chunks$c spl64" UMPC_ROLLOUT_CHUNKS=$c $B --steps-per-launch 64; run "chunks$c spl250" UMPC_ROLLOUT_CHUNKS=$c $B --steps-per-launch 250 --steps 10; done
for c in 8; do run "chunks$c 56832" UMPC_ROLLOUT_CHUNKS=$c $B --envs 56832; run "chunks$c 61568(13/SM)" UMPC_ROLLOUT_CHUNKS=$c $B --envs 61568; done
timeout 600 python -m pytest tests/test_gpu_rollout.py -m gpu -q --tb=short -x 2>&1 | tail -3 > gpurun_out/r02i_pytest.log
