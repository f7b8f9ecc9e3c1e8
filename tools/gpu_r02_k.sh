#!/bin/bash
# round 2, pass k: time-parallel tests; small-N tables per mode; launch lists
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q > gpurun_out/k_pytest_tp.log 2>&1; echo "pytest tp rc=$?"
tail -15 gpurun_out/k_pytest_tp.log
for mode in summary summary_faults full_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,32768x20000,65536x10000,131072x10000 --chunks 0,auto > gpurun_out/k_small_n_$mode.log 2>&1
  cat gpurun_out/k_small_n_$mode.log
done
for mode in summary full_faults; do
  for shape in 4096x100000 65536x10000; do
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/k_launches_${mode}_${shape}.csv python tools/small_n.py --mode $mode --shapes $shape --chunks auto > gpurun_out/k_ncu_${mode}_${shape}.log 2>&1
    python - <<P
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/k_launches_${mode}_${shape}.csv") if l.startswith('"'))]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
print("${mode} ${shape}")
for r in rows[-7:]:
    print("  ", r[ki][:70], r[vi], r[ui])
P
  done
done
