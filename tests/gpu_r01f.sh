#!/bin/bash
# round-1 re-entry check: whole GPU parity suite, headline bench (both arms), config-5 sweep, ncu of the dir-tran rollout
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short -x 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 900 python tests/dirtran_sweep.py --B 4096,65536,262144 --N 3,10,20 --steps 10 > gpurun_out/dirtran_sweep.log 2>&1
CMD="python tests/dirtran_sweep.py --B 65536 --N 10 --steps 5"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/dt_launches.csv $CMD > gpurun_out/dt_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dt_rollout -s 1 -c 1 -f -o gpurun_out/dt_prof $CMD > gpurun_out/dt_ncu_full.log 2>&1
ls -la gpurun_out
