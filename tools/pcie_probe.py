#!/usr/bin/env python
"""What the host link can do, to read bench.py's `e2e` against: every rank copies pinned host
buffers to its GPU and back at the same time (two streams), in the 8 : 5 byte ratio of the replay
(int64 depth in, int32 replicas + uint8 events out), with no kernel at all. Prints per-rank and
whole-box GB/s and the decisions/s that bandwidth alone would allow.

    python tools/pcie_probe.py                  # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29512 tools/pcie_probe.py
"""
from __future__ import annotations

import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dec = 1 << 27  # decisions per pass: 1 GiB in, 640 MiB out
    h_in = torch.empty(dec, dtype=torch.int64).pin_memory()
    h_out = torch.empty(dec * 5, dtype=torch.uint8).pin_memory()
    h_in.zero_(), h_out.zero_()
    d_in = torch.empty(dec, dtype=torch.int64, device="cuda")
    d_out = torch.zeros(dec * 5, dtype=torch.uint8, device="cuda")
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("h2d", "d2h", "both"):
        for it in range(3):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            s_in.wait_stream(torch.cuda.current_stream()), s_out.wait_stream(torch.cuda.current_stream())
            for rep in range(4):
                if mode != "d2h":
                    with torch.cuda.stream(s_in):
                        d_in.copy_(h_in, non_blocking=True)
                if mode != "h2d":
                    with torch.cuda.stream(s_out):
                        h_out.copy_(d_out, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s_in), torch.cuda.current_stream().wait_stream(s_out)
            b.record()
            torch.cuda.synchronize()
            ms = torch.tensor([a.elapsed_time(b)], device="cuda")
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms)
        nbytes = 4 * dec * ((8 if mode != "d2h" else 0) + (5 if mode != "h2d" else 0))
        res[mode] = {"ms_max_over_ranks": ms, "GBps_per_gpu": nbytes / ms / 1e6, "GBps_box": world * nbytes / ms / 1e6}
    res["decisions_per_s_allowed_by_link"] = world * 4 * dec / (res["both"]["ms_max_over_ranks"] * 1e-3)
    if rank == 0:
        print(json.dumps({"n_gpus": world, **res}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
