#!/bin/bash
# config-5 sweep, large batches (1M and 4M envs; 4M x N=20 is a 58 GB slab), and the 2-GPU stdout check
mkdir -p gpurun_out
timeout 1500 python tests/dirtran_sweep.py --B 1024,16384,1048576,4194304 --N 3,5,10,15,20 --steps 10 --max-gb 80 > gpurun_out/dirtran_sweep_big.log 2>&1
