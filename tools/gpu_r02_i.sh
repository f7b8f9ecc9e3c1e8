#!/bin/bash
# round 2, pass i: time-parallel summary / failure-mask replays: tests, then the small-N table per mode
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q > gpurun_out/i_pytest_tp.log 2>&1; echo "pytest tp rc=$?"
tail -25 gpurun_out/i_pytest_tp.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/i_pytest.log
for mode in summary summary_faults full_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,65536x10000,131072x10000,196608x8192 --chunks 0,auto > gpurun_out/i_small_n_$mode.log 2>&1
  cat gpurun_out/i_small_n_$mode.log
done
