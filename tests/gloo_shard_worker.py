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
from kube_sqs_autoscaler_b200.shard import SharedHostSegment, gather_summaries, shard_range  # noqa: E402

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
    # the same gather through a host segment every rank maps (what bench.py's config5 does with page-locked
    # slices and device-to-host copies; here plain host copies, and slices padded to whole pages)
    import mmap

    per = ((N_TOTAL + world - 1) // world * 64 + mmap.PAGESIZE - 1) // mmap.PAGESIZE * mmap.PAGESIZE
    seg = SharedHostSegment(per * world, rank * per, (rank + 1) * per)
    if seg.ok:
        mine = np.ascontiguousarray(res.summary).view(np.uint8).reshape(-1)
        seg.array[rank * per:rank * per + mine.size] = mine
        seg.barrier()
        if rank == 0:
            for r in range(world):
                rb, re_ = shard_range(N_TOTAL, world, r)
                got = bytes(seg.array[r * per:r * per + (re_ - rb) * 64])
                assert got == full[rb:re_].tobytes(), f"shared segment: slice of rank {r} differs"
        seg.close()
    unaligned = SharedHostSegment(per * world, rank * per + 8, (rank + 1) * per)  # not page-aligned: refused on every rank
    assert not unaligned.ok
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
