#!/bin/bash
# dir-tran path on the GPU: parity tests, then a first timing of the rollout kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_dirtran.py -q --tb=short -x 2>&1 | tail -25 > gpurun_out/pytest_dirtran.log
timeout 900 python tests/dirtran_sweep.py ${SWEEP_ARGS:-} > gpurun_out/dirtran_sweep.log 2>&1
