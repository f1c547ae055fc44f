#!/bin/bash
# r01j: whole GPU suite, config-5 bench lines, transient sweep, launch list + full ncu capture of dt_rollout_kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload dirtran --horizon 10 --steps 5 > gpurun_out/bench_dt_N10.json 2> gpurun_out/bench_dt_N10.err
for N in 3 5 15 20; do timeout 600 python bench.py --workload dirtran --horizon $N --steps 5 --no-cpu > gpurun_out/bench_dt_N$N.json 2>> gpurun_out/bench_dt.err; done
timeout 900 python tests/dirtran_sweep.py --B 65536,1048576 --N 3,5,10,15,20 --steps 10 > gpurun_out/dirtran_sweep.log 2>&1
CMD="python bench.py --workload dirtran --horizon 10 --steps 2 --warmup 3 --no-cpu --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/dt_launches.csv $CMD > gpurun_out/dt_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dt_rollout -s 3 -c 1 -f -o gpurun_out/dt_prof $CMD > gpurun_out/dt_ncu_full.log 2>&1
