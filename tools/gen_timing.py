#!/usr/bin/env python
"""Time the around-the-path kernels (not part of any timed step): synthetic depth generators,
failure-mask generator, trajectory fingerprint. CUDA events, best of 3 after a warm-up."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import kube_sqs_autoscaler_b200 as ksa  # noqa: E402


def best(fn, reps=3):
    ts = []
    for i in range(reps + 1):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i:
            ts.append(a.elapsed_time(b))
    return min(ts)


def main():
    eng = ksa.ReplayEngine(0)
    N, T = 1 << 20, 2000
    out = {}
    buf = torch.empty((T, 1, N), dtype=torch.int64, device=eng.dev)
    g = ksa.GenParams(seed=1)
    ms = best(lambda: eng.gen_traces(g, 0, T, 1, 0, N, N, out=buf))
    out["gen_bands"] = {"ms": ms, "GBps_written": 8 * N * T / ms / 1e6}
    gw = ksa.GenParams(seed=1, family=ksa.GEN_WALK, low_max=0, high_min=1, high_max=500, walk_step=15)
    walk = eng.gen_walk_init(gw, 0, N)
    ms = best(lambda: eng.gen_traces(gw, 0, T, 1, 0, N, N, walk=walk.clone(), out=buf))
    out["gen_walk"] = {"ms": ms, "GBps_written": 8 * N * T / ms / 1e6}
    buf3 = torch.empty((T // 2, 3, N), dtype=torch.int64, device=eng.dev)
    g3 = ksa.GenParams(seed=1, low_max=25, high_min=250, high_max=1000, overflow_log2=20)
    ms = best(lambda: eng.gen_traces(g3, 0, T // 2, 3, 0, N, N, out=buf3))
    out["gen_bands_3_attributes"] = {"ms": ms, "GBps_written": 24 * N * (T // 2) / ms / 1e6}
    fb = torch.empty((T, N), dtype=torch.uint8, device=eng.dev)
    ms = best(lambda: eng.gen_faults(1, 5, 4, 4, 0, T, 0, N, N, out=fb))
    out["gen_faults"] = {"ms": ms, "GBps_written": N * T / ms / 1e6}
    r = torch.zeros((T, N), dtype=torch.int32, device=eng.dev)
    e = torch.zeros((T, N), dtype=torch.uint8, device=eng.dev)
    ms = best(lambda: eng.traj_hash(r, e, T, N, N))
    out["traj_hash"] = {"ms": ms, "GBps_read": 5 * N * T / ms / 1e6}
    print(json.dumps({"workload": f"{N} traces x {T} ticks", **out}))


if __name__ == "__main__":
    main()
