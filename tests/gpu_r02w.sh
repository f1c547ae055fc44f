#!/bin/bash
# generic solver after the register triangular solves: parity tests + bench lines
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_gqp.py -x -m gpu -q --tb=short 2>&1 | tail -25 > $O/r02w_pytest_gqp.log
for N in 3 5 8; do timeout 900 python bench.py --workload umpc-n --horizon $N --no-cpu > $O/r02w_umpc_n$N.json 2> $O/r02w_umpc_n$N.err; done
timeout 900 python bench.py --workload quad > $O/r02w_quad.json 2> $O/r02w_quad.err
timeout 900 python bench.py --workload fused-wlqp --no-cpu > $O/r02w_fused_wlqp.json 2> $O/r02w_fused_wlqp.err
