#!/bin/bash
# ncu: launch list + full capture of the round-2 rollout kernel (after the same command ran clean)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/r02f_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02f_launches.csv $CMD > gpurun_out/r02f_ncu_launch.log 2>&1
$CMD > gpurun_out/r02f_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:umpc_rollout -s 3 -c 1 -f -o gpurun_out/r02f_prof $CMD > gpurun_out/r02f_ncu_full.log 2>&1
ls -la gpurun_out | tail -5
