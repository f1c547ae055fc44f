#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gqp.py -m gpu -q --tb=short 2>&1 | tail -40 > gpurun_out/r02k_pytest_gqp.log
