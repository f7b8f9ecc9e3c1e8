#!/bin/bash
# round 2, pass t: ksa_sweep_reduce with a cluster of CTAs per 32 traces (few traces): test, shapes
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "sweep_reduce" 2>&1 | tail -4
for sh in -1 0 101 102 103; do
  echo "KSA_REDUCE_SHAPE=$sh"
  KSA_REDUCE_SHAPE=$sh timeout 600 python tools/run_configs.py --config 4 2>&1 | grep sweep_reduce | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('  ms', d['ms'], 'GB/s', round(d['GBps'], 1), 'frac', round(d['frac_of_measured_peak'], 3), d['parity'])"
done 2>&1 | tee gpurun_out/t_reduce.log
