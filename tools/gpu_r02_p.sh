#!/bin/bash
# round 2, pass p: bench line with small_n and the overlapped config5 gather (1 GPU)
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/p_bench.err
python - <<'P'
import json
d=json.loads(open("gpurun_out/p_bench.json").read().strip().splitlines()[-1])
print(json.dumps(d["other_kernels"]["config5"], indent=1))
print(json.dumps(d["other_kernels"]["small_n"], indent=1))
print("value", d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
P
