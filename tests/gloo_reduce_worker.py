"""Worker for test_two_rank_gloo_allreduces_policy_costs: each rank scores its shard of traces under
a small policy grid (oracle replay + the numpy restatement of ksa_sweep_reduce stand in for the GPU
on this box), then shard.allreduce_policy_costs sums the [P] totals over gloo."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from kube_sqs_autoscaler_b200 import Policy  # noqa: E402
from kube_sqs_autoscaler_b200.engine import GenParams  # noqa: E402
from kube_sqs_autoscaler_b200.shard import allreduce_policy_costs, shard_range  # noqa: E402

N_TOTAL, T = 777, 64
S = 10**9
POLICIES = [Policy(attr_mask=1, scale_up_messages=60 + 30 * i, scale_down_messages=5 + 5 * i,
                   scale_up_cool_down=(i % 3) * 10 * S, scale_down_cool_down=(i % 2) * 30 * S, max_pods=3 + i)
            for i in range(6)]
GEN = GenParams(seed=0x5EED0004)
WEIGHTS = dict(w_pod_ticks=3, w_saturated=50, w_churn=7, w_blocked=-2)


def shard_costs(b: int, n: int):
    from oracle import policy_py
    from tests import oracle_lib

    depth = oracle_lib.gen_traces(GEN.to_c(), 0, T, 1, b, n, n)
    st = oracle_lib.fresh_state(GEN.seed, b, n, 1, 5)
    summ = np.stack([oracle_lib.replay(p.to_c(), depth, T, 1, n, n, 0, *st, want_full=False, want_summary=True,
                                       threads=1).summary for p in POLICIES])
    return policy_py.sweep_reduce(summ, **WEIGHTS)


def main() -> None:
    import torch
    import torch.distributed as dist

    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    b, e = shard_range(N_TOTAL, world, rank)
    cost, _, _ = shard_costs(b, e - b)
    total = allreduce_policy_costs(torch.from_numpy(cost.copy()))
    if rank == 0:
        np.save(sys.argv[1], total.numpy())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
