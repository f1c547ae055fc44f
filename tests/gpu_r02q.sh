#!/bin/bash
# converged mode on chip: bench + ncu full capture of the kernel
mkdir -p gpurun_out
O=gpurun_out
B="python bench.py --no-cpu --no-e2e --converged"
timeout 900 python -m pytest tests/test_gpu_converged.py -x -m gpu -q --tb=short 2>&1 | tail -5 > $O/r02q_pytest_conv.log
timeout 600 $B --steps 20 --warmup 3 > $O/r02q_conv_tmem.json 2> $O/r02q_conv_tmem.err
timeout 600 $B --steps 20 --warmup 3 --envs 18944 > $O/r02q_conv_tmem_18944.json 2>> $O/r02q_conv_tmem.err
CMD="$B --steps 3 --warmup 3 --envs 18944"
$CMD > $O/r02q_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:umpc_rollout -s 3 -c 1 -f -o $O/r02q_prof $CMD > $O/r02q_ncu_full.log 2>&1
