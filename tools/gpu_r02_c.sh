#!/bin/bash
# round 2, third GPU pass: parity incl. the new paths, small-N numbers, ncu of the failure-aware summary kernel
set -u
mkdir -p gpurun_out
T=kube_sqs_autoscaler_b200/libksareplay_tuning.so
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -15 gpurun_out/c_pytest.log
timeout 900 python tools/small_n.py > gpurun_out/c_small_n.log 2>&1
[ -f $T ] && KSA_LIBRARY=$PWD/$T timeout 600 python tools/run_configs.py --config 6 --sweep-tune > gpurun_out/c_faults_tune.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ksa_replay_kernel -s 5 -c 1 -o gpurun_out/c_ncu_summary_faults python tools/run_configs.py --config 6 --ticks 2000 --reps 1 > gpurun_out/c_ncu_faults.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c_launches_small_n.csv python tools/small_n.py --shapes 4096x100000 --chunks auto > gpurun_out/c_ncu_small_n.log 2>&1
tail -40 gpurun_out/c_small_n.log
