#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --workload quad --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:gqp_solve -s 6 -c 1 -f -o $O/r02z_prof $CMD > $O/r02z_ncu_full.log 2>&1
