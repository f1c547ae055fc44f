#!/bin/bash
# converged-mode bring-up: its own tests first (with a time limit), then the whole GPU suite, then a quick bench of the C path
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_converged.py -m gpu -q --tb=short 2>&1 | tail -40 > gpurun_out/pytest_conv.log
timeout 1200 python -m pytest tests -m gpu -q --tb=short -x --deselect tests/test_gpu_converged.py 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 --no-cpu 2>&1 | tail -1 > gpurun_out/bench_quick.log
