#!/bin/bash
# multi-GPU lines with the final kernel: headline (65,536 envs per GPU) and the two cascades at 32,768 envs per GPU
# usage: bash tests/gpu_r02o.sh N     (gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
PORT=29511
run() {  # tag, args...
  tag=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" \
    > gpurun_out/r02o_${tag}_n$N.json 2> gpurun_out/r02o_${tag}_n$N.err
  PORT=$((PORT+1))
}
run bench
run cascade_wlcon --workload cascade-wlcon --envs 32768
run cascade_wlqp --workload cascade-wlqp --envs 32768
run dirtran_N10 --workload dirtran --horizon 10
