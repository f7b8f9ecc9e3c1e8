#!/bin/bash
# round 2, pass w: whole suite; the small-N table per mode on three kinds of depths: independent values in
# three bands, a bounded random walk (long stretches on one side of a threshold), quiet after tick 300
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/w_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/w_pytest.log
tail -4 gpurun_out/w_pytest.log
SH=4096x100000,16384x40000,32768x20000,65536x10000,131072x10000,163840x8192
for mode in full summary full_faults summary_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto 2>&1 | tee -a gpurun_out/w_bands.log
  timeout 600 python tools/small_n.py --mode $mode --family walk --shapes $SH --chunks 0,auto 2>&1 | tee -a gpurun_out/w_walk.log
  timeout 600 python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --quiet-after 300 2>&1 | tee -a gpurun_out/w_quiet.log
done > /dev/null
grep -h "chunks=auto" gpurun_out/w_bands.log | cut -c1-150
