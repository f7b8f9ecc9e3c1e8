#!/bin/bash
# round 2, pass l: where the time-parallel path stops paying, per mode; warm-up length
set -u
mkdir -p gpurun_out
for mode in summary summary_faults full_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 49152x16000,65536x10000,98304x10000,131072x10000,196608x8192 --chunks 0,auto > gpurun_out/l_threshold_$mode.log 2>&1
  cat gpurun_out/l_threshold_$mode.log
done
for mode in summary summary_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,32768x20000 --chunks auto,37,74,148,296 --warm 32,64,96,128 > gpurun_out/l_warm_$mode.log 2>&1
  cat gpurun_out/l_warm_$mode.log
done
timeout 600 python tools/small_n.py --mode full_faults --shapes 4096x100000,16384x40000 --chunks auto --warm 32,64,96,128 > gpurun_out/l_warm_full_faults.log 2>&1
cat gpurun_out/l_warm_full_faults.log
