#!/bin/bash
# host-buffer paths: parity of the three (zero-copy / graph / direct), then the e2e leg of the bench: pull-window sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q --tb=short -k "host_buffer or batch" 2>&1 | tail -15 > gpurun_out/pytest_e2e.log
: > gpurun_out/e2e_paths.log
for rep in 1 2; do
for cfg in "1 0" "1 16" "1 32" "1 48" "1 64" "1 96" "1 128" "1 256"; do
  set -- $cfg
  echo -n "UMPC_E2E_ZEROCOPY=$1 UMPC_E2E_WINDOW=$2 : " >> gpurun_out/e2e_paths.log
  UMPC_E2E_ZEROCOPY=$1 UMPC_E2E_WINDOW=$2 timeout 300 python bench.py --steps 20 --no-cpu 2>&1 | python -c "
import sys,json
t=sys.stdin.read().strip().splitlines()
try:
    d=json.loads(t[-1]); print('device ms/step %.4f'%d['ms_per_step'],'e2e ms/step %.4f'%d['e2e']['ms_per_step'],'e2e env-steps/s %.4g'%d['e2e']['value'])
except Exception as e: print('FAILED',t[-3:])" >> gpurun_out/e2e_paths.log
done; done
