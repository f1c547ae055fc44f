#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_gqp.py -x -m gpu -q --tb=short 2>&1 | tail -25 > $O/r02A_pytest_gqp.log
for N in 3 5 8; do timeout 900 python bench.py --workload umpc-n --horizon $N --no-cpu > $O/r02A_umpc_n$N.json 2> $O/r02A_umpc_n$N.err; done
timeout 900 python bench.py --workload quad > $O/r02A_quad.json 2> $O/r02A_quad.err
timeout 900 python bench.py --workload fused-wlqp --no-cpu > $O/r02A_fused_wlqp.json 2> $O/r02A_fused_wlqp.err
CMD="python bench.py --workload umpc-n --horizon 3 --steps 2 --warmup 1 --no-cpu"
CMD="python bench.py --workload quad --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:gqp_solve -s 6 -c 1 -f -o $O/r02A_prof $CMD > $O/r02A_ncu_full.log 2>&1
