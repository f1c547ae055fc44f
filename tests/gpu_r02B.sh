#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -25 > $O/r02B_pytest_gpu.log
CMD="python bench.py --workload umpc-n --horizon 3 --steps 2 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:gqp_solve -s 4 -c 1 -f -o $O/r02B_prof $CMD > $O/r02B_ncu_full.log 2>&1
