#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; nproc >> gpurun_out/topo.txt; numactl -H >> gpurun_out/topo.txt 2>&1
G=${G:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G > gpurun_out/bench_n$G.json 2> gpurun_out/bench_n$G.err
