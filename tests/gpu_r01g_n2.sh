#!/bin/bash
# 2-GPU lines: headline workload and config 5 (N=10), both under torchrun as the driver launches them
mkdir -p gpurun_out
G=${G:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G > gpurun_out/bench_n$G.json 2> gpurun_out/bench_n$G.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $G --workload dirtran --horizon 10 --steps 5 > gpurun_out/bench_dt_N10_n$G.json 2> gpurun_out/bench_dt_N10_n$G.err
for f in gpurun_out/*.err; do tail -n 3 $f; done; true
