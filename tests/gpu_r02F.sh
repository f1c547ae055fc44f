#!/bin/bash
# final state of the round: the driver's sequence (smoke, GPU tests, default bench both arms) + launch list + full capture
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02F_smoke.log 2>&1
timeout 1500 python -m pytest tests -x -m gpu -q --tb=short 2>&1 | tail -15 > $O/r02F_pytest_gpu.log
timeout 900 python bench.py --impl reference > $O/r02F_bench_ref.json 2> $O/r02F_bench_ref.err
timeout 900 python bench.py > $O/r02F_bench_n1.json 2> $O/r02F_bench_n1.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e"
$CMD > $O/r02F_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/r02F_launches.csv $CMD > $O/r02F_ncu_launch.log 2>&1
$CMD > $O/r02F_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:umpc_rollout -s 3 -c 1 -f -o $O/r02F_prof $CMD > $O/r02F_ncu_full.log 2>&1
