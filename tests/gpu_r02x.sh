#!/bin/bash
# ncu capture of the generic warp-per-QP solver (UprightMPC2 N = 3 through umg + gqp)
mkdir -p gpurun_out
O=gpurun_out
CMD="python bench.py --workload umpc-n --horizon 3 --steps 2 --warmup 1 --no-cpu"
$CMD > $O/r02x_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gqp_solve -s 4 -c 1 -f -o $O/r02x_prof $CMD > $O/r02x_ncu_full.log 2>&1
