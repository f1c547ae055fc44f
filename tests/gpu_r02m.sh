#!/bin/bash
# full GPU suite with the tightened tolerances, default bench both arms, ncu launch list + full capture of the final rollout kernel
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02m_smoke.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -15 > gpurun_out/r02m_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r02m_bench_n1.json 2> gpurun_out/r02m_bench_n1.err
timeout 900 python bench.py --impl reference > gpurun_out/r02m_bench_ref.json 2> gpurun_out/r02m_bench_ref.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/r02m_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02m_launches.csv $CMD > gpurun_out/r02m_ncu_launch.log 2>&1
$CMD > gpurun_out/r02m_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:umpc_rollout -s 3 -c 1 -f -o gpurun_out/r02m_prof $CMD > gpurun_out/r02m_ncu_full.log 2>&1
