"""Multi-GPU layout: traces are independent (the only state is per trace: Spec.Replicas and the two
stamps, main.go:37-38 + scale/scale.go:60), so the trace axis is cut into contiguous shards, one
per rank / GPU, with no data-path collective. The only cross-rank step is a gather of the 64-byte
per-trace summaries to rank 0: through a host segment that every rank of the box maps and writes its
own slice of (SharedHostSegment: every GPU's own host link carries its shard, and the copies can run
while the next chunk is replayed), or with torch.distributed (gather_summaries; plumbing, not the
product).
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


def _cpu_list(text: str) -> set[int]:
    """'0-3,8,10-11' (sysfs cpulist) -> {0, 1, 2, 3, 8, 10, 11}."""
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def host_node_of_gpu(pci_bus_id: str, sysfs: str = "/sys") -> tuple[int, set[int]]:
    """(NUMA node, its CPUs) of the PCI device `pci_bus_id` ('0000:1b:00.0'; any case), read from sysfs;
    (-1, set()) when the platform does not say (one node, a VM without topology, no such device)."""
    import os

    try:
        with open(os.path.join(sysfs, "bus/pci/devices", pci_bus_id.lower(), "numa_node")) as f:
            node = int(f.read().strip())
        if node < 0:
            return -1, set()
        with open(os.path.join(sysfs, "devices/system/node", f"node{node}", "cpulist")) as f:
            return node, _cpu_list(f.read())
    except (OSError, ValueError):
        return -1, set()


def bind_to_gpu_node(device: int, sysfs: str = "/sys") -> dict:
    """One process per GPU: run this rank's host threads on the CPUs of the NUMA node the GPU hangs off,
    so that the page-locked buffers it allocates afterwards (first touch: ksa_host_alloc, SharedHostSegment)
    and the copies between them and the GPU stay on that node's memory and PCIe root. Call before
    allocating host buffers. Never fails: returns what it found and did
    ({"node": n, "cpus": k, "bound": bool, "why": ...}); a box that does not expose the topology, or a
    CPU set (cgroup / taskset) that shares nothing with the node, is left as it is."""
    import os

    import torch

    info = {"node": -1, "cpus": 0, "bound": False}
    try:
        prop = torch.cuda.get_device_properties(device)
        bus = f"{getattr(prop, 'pci_domain_id', 0):04x}:{prop.pci_bus_id:02x}:{prop.pci_device_id:02x}.0"
    except Exception as exc:  # (no CUDA, or a torch without the PCI fields)
        info["why"] = f"no PCI address for device {device}: {exc}"
        return info
    node, cpus = host_node_of_gpu(bus, sysfs)
    info["pci"] = bus
    info["node"] = node
    if node < 0:
        info["why"] = "the platform reports no NUMA node for the GPU"
        return info
    try:
        allowed = os.sched_getaffinity(0)
        mine = cpus & allowed
        info["cpus"] = len(mine)
        if not mine:
            info["why"] = "none of the node's CPUs is in this process's CPU set"
            return info
        if mine != allowed:
            os.sched_setaffinity(0, mine)
        info["bound"] = True
    except OSError as exc:
        info["why"] = f"sched_setaffinity: {exc}"
    return info


class SharedHostSegment:
    """`nbytes` of host memory that every rank of the box maps: a tmpfs file under /dev/shm created by
    rank 0 (torch.distributed, any backend, carries its name). This rank owns bytes [lo, hi) of it and,
    given the loaded library (`lib`: libksareplay through ctypes), page-locks them with
    ksa_host_register so that device-to-host copies into them run at the link's rate and
    asynchronously; `address(offset)` is where to copy to. After `barrier()` rank 0 reads the whole
    segment through `array` (uint8). `ok` is False, on every rank, when /dev/shm cannot hold the
    segment or the slice is not page-aligned: the caller then gathers another way."""

    def __init__(self, nbytes: int, lo: int, hi: int, lib=None, group=None):
        import mmap
        import os

        import torch.distributed as dist

        self._dist, self._group, self._lib = dist, group, lib
        self.rank = dist.get_rank(group)
        self.path = self._map = self._registered = self.array = None
        aligned = lo % mmap.PAGESIZE == 0 and hi % mmap.PAGESIZE == 0
        flags = [None] * dist.get_world_size(group)
        dist.all_gather_object(flags, bool(aligned), group=group)
        box = [None]
        if self.rank == 0 and all(flags):
            try:
                vfs = os.statvfs("/dev/shm")
                if vfs.f_bavail * vfs.f_frsize > nbytes + (256 << 20):
                    box[0] = f"/dev/shm/ksa_segment_{os.getpid()}_{os.environ.get('MASTER_PORT', '0')}"
                    with open(box[0], "wb") as f:
                        f.truncate(nbytes)
            except OSError:
                box[0] = None
        dist.broadcast_object_list(box, src=0, group=group)
        self.ok = box[0] is not None
        if not self.ok:
            return
        self.path = box[0]
        with open(self.path, "r+b") as f:
            self._map = mmap.mmap(f.fileno(), nbytes)
        self.array = np.frombuffer(self._map, dtype=np.uint8)
        self.base = self.array.ctypes.data
        self.lo, self.hi = lo, hi
        self.array[lo:hi] = 0  # touch this rank's pages before they are locked
        if lib is not None and hi > lo:
            from . import _abi

            _abi.check(lib.ksa_host_register(self.base + lo, hi - lo))
            self._registered = self.base + lo
        dist.barrier(group=group)

    def address(self, offset: int = 0) -> int:
        """Host address of byte `offset` of this rank's slice."""
        return self.base + self.lo + offset

    def barrier(self) -> None:
        """Every rank's copies have been waited for by their owners: rank 0 may read `array`."""
        self._dist.barrier(group=self._group)

    def close(self) -> None:
        if not self.ok:
            return
        if self._registered:
            self._lib.ksa_host_unregister(self._registered)
            self._registered = None
        self.array = None
        self._dist.barrier(group=self._group)
        try:
            self._map.close()
        except BufferError:  # (a view of the mapping is still alive somewhere: the file goes anyway)
            pass
        if self.rank == 0:
            import os

            os.unlink(self.path)
        self.ok = False


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
