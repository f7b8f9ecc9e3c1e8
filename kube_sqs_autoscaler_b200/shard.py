"""Multi-GPU layout: traces are independent (the only state is per trace: Spec.Replicas and the two
stamps, main.go:37-38 + scale/scale.go:60), so the trace axis is cut into contiguous shards, one
per rank / GPU, with no data-path collective. The only cross-rank step is a gather of the 64-byte
per-trace summaries to rank 0 (torch.distributed over NCCL or gloo; plumbing, not the product).
"""
from __future__ import annotations

import numpy as np


def shard_range(n_total: int, world: int, rank: int) -> tuple[int, int]:
    """[begin, end) of the traces rank `rank` of `world` owns; sizes differ by at most one and the
    concatenation over ranks is exactly range(n_total)."""
    if world < 1 or not (0 <= rank < world) or n_total < 0:
        raise ValueError(f"bad shard request n_total={n_total} world={world} rank={rank}")
    base, extra = divmod(n_total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def gather_summaries(local: np.ndarray, n_total: int, group=None):
    """Gather per-trace summary records (structured array, SUMMARY_DTYPE) from every rank to rank 0,
    ordered by global trace id. Returns the full array on rank 0, None elsewhere. Works with any
    initialised torch.distributed backend (NCCL moves through a device staging tensor)."""
    import torch
    import torch.distributed as dist

    from ._abi import SUMMARY_DTYPE

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    begin, end = shard_range(n_total, world, rank)
    assert local.dtype == SUMMARY_DTYPE and local.size == end - begin
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    max_count = max(shard_range(n_total, world, r)[1] - shard_range(n_total, world, r)[0] for r in range(world))
    send = torch.zeros(max_count * SUMMARY_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    raw = torch.from_numpy(np.ascontiguousarray(local).view(np.uint8).reshape(-1).copy())
    send[: raw.numel()] = raw.to(dev)
    bufs = [torch.empty_like(send) for _ in range(world)] if rank == 0 else None
    dist.gather(send, bufs, dst=0, group=group)
    if rank != 0:
        return None
    out = np.empty(n_total, dtype=SUMMARY_DTYPE)
    for r in range(world):
        b, e = shard_range(n_total, world, r)
        out[b:e] = bufs[r].cpu().numpy()[: (e - b) * SUMMARY_DTYPE.itemsize].view(SUMMARY_DTYPE)
    return out


def allreduce_policy_costs(policy_cost, group=None):
    """Sum the per-policy cost vectors of ksa_sweep_reduce over the ranks (in place; every rank gets
    the totals). This is the one cross-GPU step of the engine that is a collective: each rank scored
    its own shard of traces, the [P] int64 totals are all-reduced over NCCL (or gloo on CPU).
    Wrapping int64 sums are order-independent, so the result is exact."""
    import torch.distributed as dist

    dist.all_reduce(policy_cost, op=dist.ReduceOp.SUM, group=group)
    return policy_cost
