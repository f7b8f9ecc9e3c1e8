#!/bin/bash
# round 2, pass o: whole suite with the chunked summary / failure-mask paths; final small-N table per mode
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/o_pytest.log
tail -4 gpurun_out/o_pytest.log
for mode in full summary summary_faults full_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,32768x20000,49152x16000,65536x10000,81920x10000,98304x10000,131072x10000,163840x8192,196608x8192 --chunks 0,auto > gpurun_out/o_small_n_$mode.log 2>&1
  cat gpurun_out/o_small_n_$mode.log
done
