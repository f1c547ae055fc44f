#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then after the whole suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short 2>&1 | tail -15 > gpurun_out/r02n_zz_alone.log
timeout 1500 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -15 > gpurun_out/r02n_pytest_gpu.log
