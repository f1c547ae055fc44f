#!/bin/bash
mkdir -p gpurun_out
for w in quad fused-wlqp; do timeout 600 python bench.py --workload $w --steps 5 > gpurun_out/r02l_bench_$w.json 2> gpurun_out/r02l_bench_$w.err; done
for n in 3 5 8; do timeout 600 python bench.py --workload umpc-n --horizon $n --steps 5 --no-cpu > gpurun_out/r02l_bench_umpc_n$n.json 2> gpurun_out/r02l_bench_umpc_n$n.err; done
timeout 600 python bench.py --converged --no-cpu --steps 10 > gpurun_out/r02l_bench_converged.json 2> gpurun_out/r02l_bench_converged.err
timeout 600 python bench.py --workload fused-wlqp --envs 4096 --steps 5 --no-cpu > gpurun_out/r02l_bench_fused_4096.json 2>/dev/null
