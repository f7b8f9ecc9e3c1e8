#!/bin/bash
# Pass aj: replica offsets written, not added, where 8 adjacent traces have constant chunks.
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q 2>&1 | tail -3
for s in 121 122 123; do timeout 300 python tools/fuzz_timepar.py --cases 500 --seed $s | cut -c1-500 | tail -3; done
SH=4096x100000,16384x40000,65536x10000,131072x10000
python tools/small_n.py --mode full --shapes $SH --chunks auto --quiet-after 300
python tools/small_n.py --mode full_faults --shapes $SH --chunks auto --quiet-after 300
python tools/small_n.py --mode full --shapes $SH --chunks auto
python tools/small_n.py --mode full --shapes 4096x100000,16384x40000 --chunks auto --family walk
