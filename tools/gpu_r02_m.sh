#!/bin/bash
# round 2, pass m: wave-aligned chunk counts: the table per mode again
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q > gpurun_out/m_pytest_tp.log 2>&1; echo "pytest tp rc=$?"
tail -5 gpurun_out/m_pytest_tp.log
for mode in summary summary_faults full_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,32768x20000,49152x16000,65536x10000,98304x10000,131072x10000,196608x8192 --chunks 0,auto > gpurun_out/m_small_n_$mode.log 2>&1
  cat gpurun_out/m_small_n_$mode.log
done
