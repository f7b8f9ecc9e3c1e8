#!/usr/bin/env python
"""Replay at trace counts that do not fill the GPU: plain kernels against the time-parallel path
(csrc/ksa_timepar.cu), one line per (traces, ticks, chunks).

    python tools/small_n.py [--mode full|summary|full_faults|summary_faults] [--chunks 0,auto,2,4,...]

CUDA events around the call (four or five launches on the time-parallel path), best of 4 after 2
warm-ups; 13 B per decision in full mode, 8 in summary mode, one more with a failure mask, against
the measured copy peak. The outputs of every variant (trajectory fingerprint or the summary
records, and the carry state) are compared with the plain kernels'."""
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
    ap.add_argument("--mode", default="full", choices=["full", "summary", "full_faults", "summary_faults"])
    ap.add_argument("--quiet-after", type=int, default=0,
                    help="ticks after which every depth is 50 (between the thresholds: no scale event ever again, so a "
                         "chunk started from a guessed replica count never meets the true trajectory); 0 = off")
    ap.add_argument("--steady-after", type=int, default=0,
                    help="ticks after which odd traces stay at depth 500 (saturated: ScaleUp every cool-down period) and even "
                         "ones at 0 (idle: ScaleDown every period); 0 = off")
    ap.add_argument("--family", default="bands", choices=["bands", "walk"],
                    help="depth generator: independent values in three bands, or a bounded random walk (step 15, wall 500: long "
                         "stretches on one side of a threshold, and between the two)")
    ap.add_argument("--fail-log2", default="5,4,4", help="failure mask: poll, up, down bits each with probability 2^-k (0 = never)")
    ap.add_argument("--warm", default="", help="comma list of KSA_TIMEPAR_WARM values to try (default: the library's)")
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
        if args.family == "walk":
            wgen = ksa.GenParams(seed=0x5EED0004, family=ksa.GEN_WALK, high_max=500, walk_step=15)
            depth = eng.gen_traces(wgen, 0, T, 1, 0, N, N, walk=eng.gen_walk_init(wgen, 0, N))
        else:
            depth = eng.gen_traces(gen, 0, T, 1, 0, N, N)
        if args.quiet_after:
            depth[args.quiet_after:] = 50
        if args.steady_after:
            depth[args.steady_after:, :, 0::2] = 0
            depth[args.steady_after:, :, 1::2] = 500
        replicas = torch.empty((T, N), dtype=torch.int32, device=eng.dev)
        events = torch.empty((T, N), dtype=torch.uint8, device=eng.dev)
        state0 = eng.new_state(N, seed=gen.seed, lo=1, hi=5)
        summary_mode = args.mode.startswith("summary")
        fl = [int(x) for x in args.fail_log2.split(",")]
        fail = eng.gen_faults(gen.seed, fl[0], fl[1], fl[2], 0, T, 0, N, N) if args.mode.endswith("faults") else None
        bytes_per = (8 if summary_mode else 13) + (1 if fail is not None else 0)
        ref = None
        for ch, warm in [(c, w) for c in args.chunks.split(",") for w in (args.warm.split(",") if args.warm else [""])]:
            if ch == "auto":
                os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
            else:
                os.environ["KSA_TIMEPAR_CHUNKS"] = ch
            if warm:
                os.environ["KSA_TIMEPAR_WARM"] = warm
            else:
                os.environ.pop("KSA_TIMEPAR_WARM", None)
            ts = []
            launches = 0
            for rep in range(6):
                st = state0.clone()
                replicas.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                eng.launch_count()
                a.record()
                if summary_mode:
                    summ = eng.new_summary(N)
                    a.record()
                    eng.replay_summary(pol, depth, T, 1, N, N, 0, st, summary=summ, fail=fail)
                else:
                    eng.replay_full(pol, depth, T, 1, N, N, 0, st, replicas=replicas, events=events, fail=fail)
                b.record()
                torch.cuda.synchronize()
                launches = eng.launch_count()
                if rep >= 2:
                    ts.append(a.elapsed_time(b))
            h = summ if summary_mode else eng.traj_hash(replicas, events, T, N, N)
            hs = (h, st.replicas.clone(), st.last_up.clone(), st.last_down.clone())
            if ref is None:
                ref = hs
            same = all(bool(torch.equal(x, y)) for x, y in zip(hs, ref))
            ms = min(ts)
            gbs = bytes_per * N * T / (ms * 1e-3) / 1e9
            print(f"{args.mode} N={N:7d} T={T:6d} chunks={ch:>4s} warm={warm or '-':>3s} launches={launches} {ms:8.3f} ms {N * T / (ms * 1e-3):.3e} decisions/s "
                  f"{gbs:7.1f} GB/s frac={gbs / peak:5.3f} same_as_plain={same}", flush=True)
        os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
        os.environ.pop("KSA_TIMEPAR_WARM", None)
        del depth, replicas, events


if __name__ == "__main__":
    main()
