#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r02j_chunks_equal.log; : > $O
run() { local label=$1; shift
  env "$@" 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('$label','ms/envstep %.4f'%(d['ms_per_step']/d['plant_steps_per_launch']),'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'])
except Exception as e: print('$label','FAILED',t[-1][-200:])" >> $O
}
B="python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e"
run "auto spl125" $B
for c in 2 4 8 16; do run "chunks$c spl125" UMPC_ROLLOUT_CHUNKS=$c $B; done
run "auto spl250" $B --steps-per-launch 250 --steps 10
run "auto spl25" $B --steps-per-launch 25
run "auto 56832 (12/SM)" $B --envs 56832
run "auto 61568 (13/SM)" $B --envs 61568
run "auto 75776 (16/SM)" $B --envs 75776
run "auto 32768" $B --envs 32768
run "auto 262144" $B --envs 262144
run "auto 1M" $B --envs 1048576 --steps 5
run "auto cascade-wlqp" $B --workload cascade-wlqp
run "auto cascade-wlcon" $B --workload cascade-wlcon
timeout 900 python -m pytest tests/test_gpu_rollout.py tests/test_gpu_sweep.py -m gpu -q --tb=short -x 2>&1 | tail -3 > gpurun_out/r02j_pytest.log
