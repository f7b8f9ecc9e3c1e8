#!/bin/bash
# round 2, pass r: the bench line on N GPUs as the driver launches it (arg 1 = N)
set -u
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r_bench_${N}gpu.json 2> gpurun_out/r_bench_${N}gpu.err; echo "bench ${N}gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r_ref_${N}gpu.json 2> gpurun_out/r_ref_${N}gpu.err; echo "ref rc=$?"
python - <<P
import json
d=json.loads(open("gpurun_out/r_bench_${N}gpu.json").read().strip().splitlines()[-1])
print(json.dumps(d["other_kernels"]["config5"], indent=1)); print("value", d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"], d["e2e"].get("frac_of_link"))
P
tail -c 500 gpurun_out/r_bench_${N}gpu.err; tail -c 300 gpurun_out/r_ref_${N}gpu.json
