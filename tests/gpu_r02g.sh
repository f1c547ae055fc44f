#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r02g_chunks.log; : > $O
run() { local label=$1; shift
  env "$@" 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('$label','ms/envstep %.4f'%(d['ms_per_step']/d['plant_steps_per_launch']),'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'],'clk',d['clocks']['sm_mhz'],d['clocks'].get('power_w_max'))
except Exception as e: print('$label','FAILED',t[-1][-200:])" >> $O
}
B="python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rollout.py tests/test_gpu_sweep.py -m gpu -q --tb=short -x 2>&1 | tail -5 > gpurun_out/r02g_pytest.log
for c in 8 1 4 16; do run "chunks$c" UMPC_ROLLOUT_CHUNKS=$c $B; done
run "chunks8 75776" $B --envs 75776
run "chunks8 1M" $B --envs 1048576 --steps 5
run "chunks8 cascade-wlqp" $B --workload cascade-wlqp
run "chunks8 cascade-wlcon" $B --workload cascade-wlcon
run "chunks8 hl5" $B --hl-interval 5
