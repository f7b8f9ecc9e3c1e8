#!/bin/bash
# round 2, fifth GPU pass: time-parallel v2 parity and numbers (warm-up / width variations)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_timepar.py -x -q > gpurun_out/e_pytest_tp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest_tp.log
tail -25 gpurun_out/e_pytest_tp.log
timeout 600 python tools/small_n.py --chunks 0,auto > gpurun_out/e_small_n.log 2>&1
for w in 0 32 128; do KSA_TIMEPAR_WARM=$w timeout 300 python tools/small_n.py --shapes 4096x100000,65536x10000 --chunks auto > gpurun_out/e_small_n_warm$w.log 2>&1; done
KSA_TIMEPAR_WIDE=0 timeout 300 python tools/small_n.py --shapes 4096x100000,65536x10000 --chunks auto > gpurun_out/e_small_n_narrow.log 2>&1
timeout 300 python tools/small_n.py --shapes 4096x100000 --chunks 32,64,96,128,192 > gpurun_out/e_small_n_k4096.log 2>&1
timeout 300 python tools/small_n.py --shapes 65536x10000 --chunks 4,6,8,12 > gpurun_out/e_small_n_k65536.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/e_launches_small_n.csv python tools/small_n.py --shapes 4096x100000,65536x10000 --chunks auto > gpurun_out/e_ncu_small_n.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest.log
tail -5 gpurun_out/e_pytest.log
cat gpurun_out/e_small_n.log; grep auto gpurun_out/e_small_n_warm*.log gpurun_out/e_small_n_narrow.log; cat gpurun_out/e_small_n_k4096.log gpurun_out/e_small_n_k65536.log
