#!/bin/bash
# dir-tran path on the GPU: parity tests, then the B x N sweep of the rollout kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_dirtran.py -q --tb=short 2>&1 | tail -25 > gpurun_out/pytest_dirtran.log
timeout 900 python tests/dirtran_sweep.py ${SWEEP_ARGS:---B 4096,65536,262144 --N 3,10,20 --steps 10} > gpurun_out/dirtran_sweep.log 2>&1
