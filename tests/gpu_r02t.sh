#!/bin/bash
# usage: bash tests/gpu_r02t.sh N
# N=1: sustained headline run with the clock sampler's full record + default bench (new CPU sample) + host-link ceiling
# N>1: host-link ceiling from all ranks at once, next to the e2e line of the same box
N=${1:-1}
mkdir -p gpurun_out
O=gpurun_out
if [ "$N" = "1" ]; then
  BENCH_CLOCK_RECORD=$O/r02t_clock_record_steps100.csv timeout 900 python bench.py --steps 100 --warmup 3 --no-cpu --no-e2e > $O/r02t_bench_sustained_steps100.json 2> $O/r02t_sustained.err
  timeout 900 python bench.py > $O/r02t_bench_n1.json 2> $O/r02t_bench_n1.err
  timeout 600 python bench.py --workload hostlink --steps 20 > $O/r02t_hostlink_n1.json 2> $O/r02t_hostlink_n1.err
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
  timeout 600 $TR --master-port 29611 bench.py --gpus $N --workload hostlink --steps 20 > $O/r02t_hostlink_n$N.json 2> $O/r02t_hostlink_n$N.err
  timeout 900 $TR --master-port 29612 bench.py --gpus $N --no-cpu > $O/r02t_bench_n$N.json 2> $O/r02t_bench_n$N.err
  nvidia-smi topo -m > $O/r02t_topo_n$N.txt 2>&1
fi
