#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02e_smoke.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -15 > gpurun_out/r02e_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r02e_bench_n1.json 2> gpurun_out/r02e_bench_n1.err
timeout 900 python bench.py --impl reference > gpurun_out/r02e_bench_ref.json 2> gpurun_out/r02e_bench_ref.err
timeout 600 python bench.py --workload c1 --steps 3 > gpurun_out/r02e_bench_c1.json 2> gpurun_out/r02e_bench_c1.err
timeout 600 python bench.py --workload qp4096 > gpurun_out/r02e_bench_qp4096.json 2> gpurun_out/r02e_bench_qp4096.err
timeout 600 python bench.py --envs 75776 --no-cpu --no-e2e > gpurun_out/r02e_bench_75776.json 2>/dev/null
timeout 600 python bench.py --workload dirtran --horizon 10 --steps 5 --no-cpu > gpurun_out/r02e_bench_dirtran_N10.json 2> gpurun_out/r02e_bench_dirtran.err
