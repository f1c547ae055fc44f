#!/bin/bash
# converged mode: persistent grid with dynamic tiles, 6 factor chunks in shared memory (default) vs all in tensor memory (variant)
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -15 > $O/r02s_pytest_gpu.log
B="python bench.py --no-cpu --no-e2e --converged --steps 20 --warmup 3"
timeout 600 $B > $O/r02s_conv_m0aaa.json 2> $O/r02s_conv.err
timeout 600 $B --envs 75776 > $O/r02s_conv_m0aaa_75776.json 2>> $O/r02s_conv.err
UMPC_B200_LIB=$PWD/build_variants/libumpc_b200_m0.so timeout 600 $B > $O/r02s_conv_m0.json 2>> $O/r02s_conv.err
UMPC_B200_LIB=$PWD/build_variants/libumpc_b200_m0.so timeout 600 $B --envs 75776 > $O/r02s_conv_m0_75776.json 2>> $O/r02s_conv.err
timeout 600 python bench.py --no-cpu --converged --steps 20 --warmup 3 > $O/r02s_conv_with_e2e.json 2>> $O/r02s_conv.err
CMD="python bench.py --no-cpu --no-e2e --converged --steps 3 --warmup 3"
$CMD > $O/r02s_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:umpc_rollout -s 3 -c 1 -f -o $O/r02s_prof $CMD > $O/r02s_ncu_full.log 2>&1
