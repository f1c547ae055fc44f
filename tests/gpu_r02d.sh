#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/r02d_ab.log; : > $O
run() { local label=$1; shift
  env "$@" 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('$label','ms/envstep %.4f'%(d['ms_per_step']/d['plant_steps_per_launch']),'value %.4g'%d['value'],'frac %.3f'%d['roofline']['frac'],'clk',d['clocks']['sm_mhz'],d['clocks'].get('power_w_max'))
except Exception as e: print('$label','FAILED',t[-1][-200:])" >> $O
}
B="python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e"
OLD=$PWD/build_variants/libumpc_b200_old.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rollout.py -m gpu -q --tb=short -x 2>&1 | tail -5 > gpurun_out/r02d_pytest.log
for spl in 1 4 16 64 125; do
  run "new spl$spl" $B --steps-per-launch $spl
  run "new spl$spl info" UMPC_ROLLOUT_INFO=1 $B --steps-per-launch $spl
  run "old spl$spl" UMPC_B200_LIB=$OLD $B --steps-per-launch $spl
done
run "new spl16 75776" $B --steps-per-launch 16 --envs 75776
run "new spl16 stagger40" UMPC_STAGGER_US=40 $B --steps-per-launch 16
