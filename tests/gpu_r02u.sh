#!/bin/bash
# row f2: flapping-wing plant tests + sweep bench lines; whole GPU suite
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_flapping.py -x -m gpu -q --tb=short 2>&1 | tail -25 > $O/r02u_pytest_flapping.log
timeout 900 python bench.py --workload flapsweep --steps 5 > $O/r02u_flapsweep_4096.json 2> $O/r02u_flapsweep.err
timeout 900 python bench.py --workload flapsweep --steps 5 --envs 262144 --no-cpu > $O/r02u_flapsweep_234256.json 2>> $O/r02u_flapsweep.err
timeout 900 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -15 > $O/r02u_pytest_gpu.log
