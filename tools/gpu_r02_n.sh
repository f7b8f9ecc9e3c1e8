#!/bin/bash
# round 2, pass n: ksa_replay_full_faults between 65536 and 262144 traces: one trace per thread (variant 9),
# four (variant 8), chunks of the time axis
set -u
mkdir -p gpurun_out
for v in 9 8; do
  echo "KSA_FAULTS_VARIANT=$v"
  KSA_FAULTS_VARIANT=$v timeout 600 python tools/small_n.py --mode full_faults --shapes 81920x10000,98304x10000,114688x10000,131072x10000,163840x8192,196608x8192,262144x8192 --chunks 0 2>&1 | tee -a gpurun_out/n_full_faults_plain.log
done
timeout 600 python tools/small_n.py --mode full_faults --shapes 81920x10000,114688x10000,163840x8192 --chunks auto,2 2>&1 | tee gpurun_out/n_full_faults_tp.log
