#!/bin/bash
# round 2, fourth GPU pass: parity, small-N after the fix-up changes, deeper-ring failure-aware variants
set -u
mkdir -p gpurun_out
T=kube_sqs_autoscaler_b200/libksareplay_tuning.so
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -15 gpurun_out/d_pytest.log
timeout 900 python tools/small_n.py --chunks 0,auto,16,32,64,96,128 > gpurun_out/d_small_n.log 2>&1
[ -f $T ] && KSA_LIBRARY=$PWD/$T timeout 600 python tools/run_configs.py --config 6 --sweep-tune > gpurun_out/d_faults_tune.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/d_launches_small_n.csv python tools/small_n.py --shapes 4096x100000,65536x10000 --chunks auto > gpurun_out/d_ncu_small_n.log 2>&1
cat gpurun_out/d_small_n.log
grep summary gpurun_out/d_faults_tune.log | sort -t: -k2 -n | head -30
