"""Worker for test_two_rank_gloo_shards_and_gathers: exercises the N>1 host path (shard_range,
global trace ids in the generator, gather_summaries) over gloo on CPU. The per-shard replay is done
by the oracle here because this box has no GPU; bench.py runs the same sharding with the CUDA
library on real GPUs."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from kube_sqs_autoscaler_b200 import Policy  # noqa: E402
from kube_sqs_autoscaler_b200.engine import GenParams  # noqa: E402
from kube_sqs_autoscaler_b200.shard import gather_summaries, shard_range  # noqa: E402

N_TOTAL, T = 1001, 96
POLICY = Policy(attr_mask=1)
GEN = GenParams(seed=0x5EED0001)


def main() -> None:
    import torch.distributed as dist

    from tests import oracle_lib

    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    b, e = shard_range(N_TOTAL, world, rank)
    n = e - b
    depth = oracle_lib.gen_traces(GEN.to_c(), 0, T, 1, b, n, n)  # global trace ids b..e-1
    res = oracle_lib.replay(POLICY.to_c(), depth, T, 1, n, n, 0,
                            *oracle_lib.fresh_state(GEN.seed, b, n, 1, 5), want_full=False,
                            want_summary=True, threads=1)
    full = gather_summaries(res.summary, N_TOTAL)
    if rank == 0:
        np.save(sys.argv[1], full)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
