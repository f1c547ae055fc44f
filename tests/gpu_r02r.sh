#!/bin/bash
# converged mode on chip with float Ruiz + fast reciprocal; the round's remaining profile lines (1 GPU)
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -15 > $O/r02r_pytest_gpu.log
B="python bench.py --no-cpu --no-e2e"
timeout 600 $B --converged --steps 20 --warmup 3 > $O/r02r_conv_tmem.json 2> $O/r02r_conv.err
timeout 600 $B --converged --steps 20 --warmup 3 --envs 18944 > $O/r02r_conv_tmem_18944.json 2>> $O/r02r_conv.err
timeout 600 $B --converged --steps 20 --warmup 3 --envs 75776 > $O/r02r_conv_tmem_75776.json 2>> $O/r02r_conv.err
timeout 900 python bench.py --workload c1 > $O/r02r_c1.json 2> $O/r02r_c1.err
timeout 900 python bench.py --workload qp4096 > $O/r02r_qp4096.json 2> $O/r02r_qp4096.err
timeout 900 python bench.py --workload quad > $O/r02r_quad.json 2> $O/r02r_quad.err
timeout 900 python bench.py --workload fused-wlqp > $O/r02r_fused_wlqp.json 2> $O/r02r_fused_wlqp.err
for N in 3 5 8; do timeout 900 python bench.py --workload umpc-n --horizon $N > $O/r02r_umpc_n$N.json 2> $O/r02r_umpc_n$N.err; done
timeout 900 $B --workload cascade-wlcon > $O/r02r_cascade_wlcon.json 2> $O/r02r_cascade_wlcon.err
timeout 900 $B --workload cascade-wlqp > $O/r02r_cascade_wlqp.json 2> $O/r02r_cascade_wlqp.err
timeout 900 $B --envs 75776 > $O/r02r_bench_75776.json 2> $O/r02r_bench_75776.err
timeout 900 $B --envs 56832 > $O/r02r_bench_56832.json 2> $O/r02r_bench_56832.err
