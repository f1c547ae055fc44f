#!/bin/bash
# converged mode on chip (tensor-memory scratch + shared-memory loop vectors) vs the streamed store; the round's profile lines
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_converged.py tests/test_gpu_parity.py -x -m gpu -q --tb=short 2>&1 | tail -15 > $O/r02p_pytest_conv.log
B="python bench.py --no-cpu --no-e2e"
timeout 600 $B --converged --steps 20 --warmup 3 > $O/r02p_conv_tmem.json 2> $O/r02p_conv_tmem.err
UMPC_CONV_STORE=glob timeout 600 $B --converged --steps 20 --warmup 3 > $O/r02p_conv_glob.json 2> $O/r02p_conv_glob.err
timeout 600 $B --converged --steps 20 --warmup 3 --envs 18944 > $O/r02p_conv_tmem_18944.json 2>> $O/r02p_conv_tmem.err
timeout 600 $B --converged --steps 20 --warmup 3 --envs 75776 > $O/r02p_conv_tmem_75776.json 2>> $O/r02p_conv_tmem.err
timeout 900 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -15 > $O/r02p_pytest_gpu.log
