#!/usr/bin/env python
"""Policy sweep + scoring on N GPUs: each rank sweeps its shard of the traces,
scores it with ksa_sweep_reduce, and the [P] totals are all-reduced over NCCL
(shard.allreduce_policy_costs). Rank 0 then repeats the whole grid alone and checks that the
all-reduced totals are identical.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/sweep_multi_gpu.py [--policies P] [--traces N] [--ticks T]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import kube_sqs_autoscaler_b200 as ksa  # noqa: E402
from kube_sqs_autoscaler_b200.engine import policy_array  # noqa: E402
from kube_sqs_autoscaler_b200.shard import allreduce_policy_costs, shard_range  # noqa: E402
from tools.run_configs import sweep_grid  # noqa: E402

W = dict(w_pod_ticks=1, w_saturated=25, w_churn=5, w_blocked=1)


def score(eng, parr, P, gen, T, b, n):
    walk = eng.gen_walk_init(gen, b, n)
    depth = eng.gen_traces(gen, 0, T, 1, b, n, n, walk=walk)
    r0 = eng.new_state(n, seed=gen.seed, lo=1, hi=5, trace0=b).replicas
    summ = torch.empty((P, n, 64), dtype=torch.uint8, device=eng.dev)
    eng.replay_sweep(parr, depth, T, 1, n, n, r0, summary=summ)
    return eng.sweep_reduce(summ, P, n, **W)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--policies", type=int, default=4096)
    ap.add_argument("--traces", type=int, default=8192)
    ap.add_argument("--ticks", type=int, default=1024)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = ksa.ReplayEngine(local)
    P, N, T = args.policies, args.traces, args.ticks
    parr = policy_array(sweep_grid()[:P])
    gen = ksa.GenParams(seed=0x5EED0004, family=ksa.GEN_WALK, high_max=500, walk_step=15)
    b, e = shard_range(N, world, rank)
    score(eng, parr, P, gen, T, b, e - b)  # warm-up
    dist.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    cost, best_policy, _ = score(eng, parr, P, gen, T, b, e - b)
    allreduce_policy_costs(cost)
    t1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([t0.elapsed_time(t1)], device=eng.dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        whole, _, _ = score(eng, parr, P, gen, T, 0, N)
        print(json.dumps({"workload": f"sweep+score {P} policies x {N} traces x {T} ticks on {world} GPUs",
                          "ms_max_over_ranks": float(ms), "decisions_per_s": P * N * T / (float(ms) * 1e-3),
                          "best_policy": int(cost.argmin()), "allreduce_bytes": 8 * P,
                          "parity": "all-reduced totals == one-GPU totals: " + ("ok" if torch.equal(whole, cost) else "MISMATCH")}),
              flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
