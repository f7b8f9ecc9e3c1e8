#!/bin/bash
# Pass al: where the chunked full replay stops paying on the four kinds of depths (49152 traces with a mask, 98304 without).
cd "$(dirname "$0")/.."
for fam in "" "--family walk" "--quiet-after 300" "--steady-after 300"; do
  python tools/small_n.py --mode full_faults --shapes 49152x16384 --chunks 0,auto $fam
  python tools/small_n.py --mode full --shapes 98304x10000 --chunks 0,auto $fam
done
