#!/bin/bash
# round 2, pass q: whole suite; traces that go quiet (the state never forgets: no guess meets it), per mode
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/q_pytest.log
tail -4 gpurun_out/q_pytest.log
for mode in full summary full_faults summary_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,65536x10000 --chunks 0,auto --quiet-after 300 2>&1 | tee -a gpurun_out/q_quiet.log
done
for mode in full summary full_faults summary_faults; do
  timeout 600 python tools/small_n.py --mode $mode --shapes 4096x100000,16384x40000,65536x10000 --chunks auto 2>&1 | tee -a gpurun_out/q_normal.log
done
