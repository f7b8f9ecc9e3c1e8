#!/bin/bash
# round 2, second GPU pass: parity, the bench line, faults / sweep numbers, two ncu captures
set -u
mkdir -p gpurun_out
T=kube_sqs_autoscaler_b200/libksareplay_tuning.so
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -3 gpurun_out/b_pytest.log
timeout 900 python bench.py > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench rc=$?"
[ -f $T ] && KSA_LIBRARY=$PWD/$T timeout 600 python tools/run_configs.py --config 6 --sweep-tune > gpurun_out/b_faults_tune.log 2>&1
timeout 300 python tools/run_configs.py --config 6 > gpurun_out/b_faults_product.log 2>&1
timeout 300 python tools/run_configs.py --config 4 > gpurun_out/b_sweep_product.log 2>&1
timeout 300 python tools/run_configs.py --config 4 --ordered-grid > gpurun_out/b_sweep_product_ordered.log 2>&1
timeout 600 python tools/sweep_variants.py --ticks 2000 > gpurun_out/b_product_variants.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ksa_replay_kernel -s 6 -c 1 -o gpurun_out/b_ncu_summary_faults python tools/run_configs.py --config 6 --ticks 2000 --reps 1 > gpurun_out/b_ncu_faults.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ksa_sweep_kernel -s 2 -c 3 -o gpurun_out/b_ncu_sweep python tools/run_configs.py --config 4 --reps 1 > gpurun_out/b_ncu_sweep.log 2>&1
tail -2 gpurun_out/b_faults_product.log
tail -c 600 gpurun_out/b_bench.err
