#!/bin/bash
# round 2, first GPU pass: parity, then tuning measurements, then one ncu capture
set -u
mkdir -p gpurun_out
T=kube_sqs_autoscaler_b200/libksareplay_tuning.so
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
KSA_LIBRARY=$PWD/$T timeout 600 python tools/run_configs.py --config 6 --sweep-tune > gpurun_out/a_faults_tune.log 2>&1
timeout 300 python tools/run_configs.py --config 6 > gpurun_out/a_faults_product.log 2>&1
KSA_LIBRARY=$PWD/$T timeout 900 python tools/run_configs.py --config 4 --sweep-tune > gpurun_out/a_sweep_tune.log 2>&1
timeout 300 python tools/run_configs.py --config 4 > gpurun_out/a_sweep_product.log 2>&1
timeout 300 python tools/run_configs.py --config 4 --ordered-grid > gpurun_out/a_sweep_product_ordered.log 2>&1
KSA_LIBRARY=$PWD/$T timeout 600 python tools/sweep_variants.py --full none --ticks 2000 > gpurun_out/a_summary_variants.log 2>&1
timeout 600 python tools/sweep_variants.py --ticks 2000 > gpurun_out/a_product_variants.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ksa_replay_kernel -c 2 -o gpurun_out/a_faults_summary python tools/run_configs.py --config 6 --ticks 1000 --reps 0 > gpurun_out/a_ncu_faults.log 2>&1
tail -3 gpurun_out/a_faults_product.log gpurun_out/a_sweep_product.log
