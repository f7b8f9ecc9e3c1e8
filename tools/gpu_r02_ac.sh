#!/bin/bash
# Pass ac: the chunks' kinds noted by pass 1 (no classify launch): tests, fuzz, timings on the four kinds of depths.
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q 2>&1 | tail -3
for s in 111 112; do timeout 300 python tools/fuzz_timepar.py --cases 500 --seed $s | cut -c1-500 | tail -3; done
SH=4096x100000,16384x40000,65536x10000
for mode in full; do
  python tools/small_n.py --mode $mode --shapes $SH --chunks auto
  python tools/small_n.py --mode $mode --family walk --shapes $SH --chunks auto
  python tools/small_n.py --mode $mode --shapes $SH --chunks auto --quiet-after 300
  python tools/small_n.py --mode $mode --shapes $SH --chunks auto --steady-after 300
done
