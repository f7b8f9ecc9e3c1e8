#!/bin/bash
# round 2, pass s: ncu --set full on the kernels of the chunked summary replay (4096 x 10^5) and on pass 1 of the
# chunked full replay with a failure mask
set -u
mkdir -p gpurun_out
timeout 600 python tools/small_n.py --mode summary --shapes 4096x100000 --chunks auto > gpurun_out/s_plain.log 2>&1 || exit 1
for k in tp_sum_warm_kernel ksa_replay_kernel tp_sum_fix_kernel tp_sum_chain_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/s_ncu_$k python tools/small_n.py --mode summary --shapes 4096x100000 --chunks auto > gpurun_out/s_ncu_$k.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tp_spec_kernel -s 2 -c 1 -f -o gpurun_out/s_ncu_tp_spec_faults python tools/small_n.py --mode full_faults --shapes 4096x100000 --chunks auto > gpurun_out/s_ncu_spec.log 2>&1
ls -la gpurun_out/*.ncu-rep
