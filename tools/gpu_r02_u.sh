#!/bin/bash
# round 2, pass u: summary replay with a failure mask: masks with and without failing scale calls
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_faults.py tests/test_ingest.py -x -q 2>&1 | tail -4
for fl in 5,4,4 5,0,0 10,0,0 0,0,0 5,12,12; do
  echo "mask $fl"
  timeout 600 python tools/small_n.py --mode summary_faults --shapes 1048576x4000 --chunks 0 --fail-log2 $fl
done 2>&1 | tee gpurun_out/u_polls.log
