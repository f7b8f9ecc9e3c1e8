#!/bin/bash
# round 2, pass j: time-parallel tests again; launch lists of the small-N runs per mode
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_timepar.py -x -q > gpurun_out/j_pytest_tp.log 2>&1; echo "pytest tp rc=$?"
tail -5 gpurun_out/j_pytest_tp.log
for mode in summary full_faults; do
  for shape in 4096x100000 65536x10000; do
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/j_launches_${mode}_${shape}.csv python tools/small_n.py --mode $mode --shapes $shape --chunks auto > gpurun_out/j_ncu_${mode}_${shape}.log 2>&1
    python - <<P
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/j_launches_${mode}_${shape}.csv") if l.startswith('"'))]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
print("${mode} ${shape}")
for r in rows[-12:]:
    print("  ", r[ki][:70], r[vi], r[ui])
P
  done
done
