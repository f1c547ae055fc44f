#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_converged.py -m gpu -q --tb=short 2>&1 | tail -30 > gpurun_out/pytest_conv.log
for tpb in 64 128; do
UMPC_CONV_TPB=$tpb timeout 600 python bench.py --converged --envs 16384 --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 > gpurun_out/bench_conv_$tpb.log
done
UMPC_CONV_TPB=128 timeout 600 python bench.py --converged --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 > gpurun_out/bench_conv_64k.log
