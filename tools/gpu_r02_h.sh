#!/bin/bash
# round 2, pass h: the whole GPU suite, smoke, both bench arms on the rebuilt tree
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/h_pytest.log
tail -4 gpurun_out/h_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/h_smoke.log 2>&1; tail -2 gpurun_out/h_smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/h_ref.json 2> gpurun_out/h_ref.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/h_ref.json; tail -c 300 gpurun_out/h_bench.err; head -c 600 gpurun_out/h_bench.json
