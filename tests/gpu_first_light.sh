#!/bin/bash
# GPU session: tests, smoke, bench (both arms), tpb sweep
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -30 > gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
for t in 128 256 32; do python bench.py --steps 10 --warmup 3 --tpb $t --no-cpu --no-e2e 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('tpb',$t,'value %.4g'%d['value'],'ms/step %.3f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'], 'peak',d['roofline']['peak'])"; done > gpurun_out/tpb_sweep.log 2>&1
nproc > gpurun_out/host.log; lscpu | head -20 >> gpurun_out/host.log
