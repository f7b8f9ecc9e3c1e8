#!/usr/bin/env python
"""Full-trajectory replay at trace counts that do not fill the GPU: plain kernels against the
time-parallel path (csrc/ksa_timepar.cu), one line per (traces, ticks, chunks).

    python tools/small_n.py [--chunks 0,auto,2,4,...]

CUDA events around the call (three launches on the time-parallel path), best of 4 after 2 warm-ups;
13 B per decision against the measured copy peak. The trajectory fingerprint of every variant is
compared with the plain kernels'."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import kube_sqs_autoscaler_b200 as ksa  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="4096x100000,16384x40000,65536x10000,98304x10000,131072x10000")
    ap.add_argument("--chunks", default="0,auto,2,4,8,16,32,64,128")
    args = ap.parse_args()
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    eng = ksa.ReplayEngine(0)
    pol = ksa.Policy(attr_mask=1)
    gen = ksa.GenParams(seed=0x5EED0001)
    for shape in args.shapes.split(","):
        N, T = (int(x) for x in shape.split("x"))
        depth = eng.gen_traces(gen, 0, T, 1, 0, N, N)
        replicas = torch.empty((T, N), dtype=torch.int32, device=eng.dev)
        events = torch.empty((T, N), dtype=torch.uint8, device=eng.dev)
        state0 = eng.new_state(N, seed=gen.seed, lo=1, hi=5)
        ref = None
        for ch in args.chunks.split(","):
            if ch == "auto":
                os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
            else:
                os.environ["KSA_TIMEPAR_CHUNKS"] = ch
            ts = []
            launches = 0
            for rep in range(6):
                st = state0.clone()
                replicas.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                eng.launch_count()
                a.record()
                eng.replay_full(pol, depth, T, 1, N, N, 0, st, replicas=replicas, events=events)
                b.record()
                torch.cuda.synchronize()
                launches = eng.launch_count()
                if rep >= 2:
                    ts.append(a.elapsed_time(b))
            h = eng.traj_hash(replicas, events, T, N, N)
            hs = (h, st.replicas.clone(), st.last_up.clone(), st.last_down.clone())
            if ref is None:
                ref = hs
            same = all(bool(torch.equal(x, y)) for x, y in zip(hs, ref))
            ms = min(ts)
            gbs = 13 * N * T / (ms * 1e-3) / 1e9
            print(f"N={N:7d} T={T:6d} chunks={ch:>4s} launches={launches} {ms:8.3f} ms {N * T / (ms * 1e-3):.3e} decisions/s "
                  f"{gbs:7.1f} GB/s frac={gbs / peak:5.3f} same_as_plain={same}", flush=True)
        os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
        del depth, replicas, events


if __name__ == "__main__":
    main()
