#!/bin/bash
# round 2, pass am: ncu --set full on the kernels that write chunks from closed forms / constant counts (full mode,
# 4096 x 10^5), and on pass 1 with the chunk kinds noted on the way. Each program ran to the end without ncu before (pass x).
cd "$(dirname "$0")/.."
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tp_fill_kernel -s 2 -c 1 -f -o gpurun_out/am_ncu_tp_fill python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto --steady-after 300 > gpurun_out/am_1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tp_shift_fill_kernel -s 2 -c 1 -f -o gpurun_out/am_ncu_tp_shift_fill python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto --quiet-after 300 > gpurun_out/am_2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tp_spec_kernel -s 4 -c 1 -f -o gpurun_out/am_ncu_tp_spec python tools/small_n.py --mode full --shapes 4096x100000 --chunks auto > gpurun_out/am_3.log 2>&1
ls -la gpurun_out/am_*.ncu-rep
