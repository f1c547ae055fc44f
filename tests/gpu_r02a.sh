#!/bin/bash
# round 2, first GPU call: whole GPU suite on the fixed library + baseline numbers of the round-1 kernel (phases, warps/SM)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short -x 2>&1 | tail -30 > gpurun_out/r02a_pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench_n1.json 2> gpurun_out/r02a_bench_n1.err
bash tests/gpu_occ.sh;  mv gpurun_out/occ_sweep.log gpurun_out/r02a_occ_sweep.log
bash tests/gpu_phase.sh; mv gpurun_out/phase_sweep.log gpurun_out/r02a_phase_sweep.log
