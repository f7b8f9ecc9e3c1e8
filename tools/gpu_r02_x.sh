#!/bin/bash
# round 2, pass x: whole suite; the small-N tables on four kinds of depths with the final library; bench line
set -u
mkdir -p gpurun_out
rm -f gpurun_out/x_*.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/x_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/x_pytest.log
tail -4 gpurun_out/x_pytest.log
SH=4096x100000,16384x40000,32768x20000,49152x16384,65536x10000,131072x10000,163840x8192
for mode in full summary full_faults summary_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto >> gpurun_out/x_bands.log 2>&1
  timeout 600 python tools/small_n.py --mode $mode --family walk --shapes $SH --chunks 0,auto >> gpurun_out/x_walk.log 2>&1
  timeout 600 python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --quiet-after 300 >> gpurun_out/x_quiet.log 2>&1
  timeout 600 python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --steady-after 300 >> gpurun_out/x_steady.log 2>&1
done
timeout 900 python bench.py > gpurun_out/x_bench.json 2> gpurun_out/x_bench.err; echo "bench rc=$?"
grep -c same_as_plain=True gpurun_out/x_*.log; grep -c same_as_plain=False gpurun_out/x_*.log
