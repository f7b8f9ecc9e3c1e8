#!/bin/bash
# Pass z: the final combine of summary mode by all lanes when every chunk fits as it is or shifted.
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q 2>&1 | tail -3
for s in 92 93 94 95; do timeout 300 python tools/fuzz_timepar.py --cases 500 --seed $s | tail -2; done
SH=4096x100000,16384x40000
for mode in summary summary_faults; do
  python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto
  python tools/small_n.py --mode $mode --family walk --shapes $SH --chunks 0,auto
  python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --quiet-after 300
  python tools/small_n.py --mode $mode --shapes $SH --chunks 0,auto --steady-after 300
done
