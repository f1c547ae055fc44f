#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_converged.py tests/test_gpu_rollout.py -x -m gpu -q --tb=short 2>&1 | tail -8 > $O/r02C_pytest_conv.log
B="python bench.py --no-cpu --no-e2e --converged --warmup 3"
timeout 600 $B --steps 10 > $O/r02C_conv_spl25.json 2> $O/r02C_conv.err
UMPC_CONV_CHUNKS=1 timeout 600 $B --steps 10 > $O/r02C_conv_spl25_nch1.json 2>> $O/r02C_conv.err
UMPC_CONV_CHUNKS=25 timeout 600 $B --steps 10 > $O/r02C_conv_spl25_nch25.json 2>> $O/r02C_conv.err
timeout 600 $B --steps 20 --steps-per-launch 1 > $O/r02C_conv_spl1.json 2>> $O/r02C_conv.err
timeout 600 $B --steps 5 --steps-per-launch 100 > $O/r02C_conv_spl100.json 2>> $O/r02C_conv.err
