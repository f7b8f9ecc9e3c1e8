#!/bin/bash
# round 2, sixth GPU pass: everything once more with the final defaults; bench line; ncu evidence
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -4 gpurun_out/f_pytest.log
timeout 900 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"
timeout 600 python tools/small_n.py --shapes 4096x100000,16384x40000,65536x10000,98304x10000,131072x10000,196608x8192 --chunks 0,auto > gpurun_out/f_small_n.log 2>&1
timeout 300 python tools/run_configs.py --config 6 > gpurun_out/f_faults_product.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ksa_replay_kernel -s 5 -c 1 -o gpurun_out/f_ncu_summary_faults python tools/run_configs.py --config 6 --ticks 2000 --reps 1 > gpurun_out/f_ncu_faults.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/f_ncu_bench.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:ksa_replay_kernel -s 3 -c 1 --csv --log-file gpurun_out/f_traffic_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/f_ncu_traffic.log 2>&1
cat gpurun_out/f_small_n.log; tail -1 gpurun_out/f_faults_product.log | cut -c1-700; tail -c 400 gpurun_out/f_bench.err
