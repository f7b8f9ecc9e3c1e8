#!/usr/bin/env python
"""Read-only, write-only and copy bandwidth of the GPU's HBM with library kernels (torch), to read the
roofline fractions of kernels whose traffic is not half reads / half writes against: MEASURED_PEAKS.json
holds the copy figure only. CUDA events, best of 5 after 2 warm-ups, 8 GiB buffers (>> L2)."""
import json

import torch


def best(fn, reps=5):
    ts = []
    for i in range(reps + 2):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(a.elapsed_time(b))
    return min(ts)


def main():
    n = 1 << 30  # int64 elements: 8 GiB
    x = torch.ones(n, dtype=torch.int64, device="cuda")
    y = torch.empty_like(x)
    nbytes = 8 * n
    out = {
        "write_only_GBps (tensor.fill_)": nbytes / best(lambda: y.fill_(7)) / 1e6,
        "read_only_GBps (tensor.sum)": nbytes / best(lambda: x.sum()) / 1e6,
        "copy_GBps (read+write bytes, y.copy_(x))": 2 * nbytes / best(lambda: y.copy_(x)) / 1e6,
    }
    x32 = torch.ones(n, dtype=torch.int32, device="cuda")
    out["widen_GBps (int32 -> int64 copy, 4 B read + 8 B written)"] = 12 * n / best(lambda: y.copy_(x32)) / 1e6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
