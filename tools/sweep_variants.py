#!/usr/bin/env python
"""Time every kernel variant of ksa_replay_full / ksa_replay_summary on one GPU.

Usage (on the GPU box): python tools/sweep_variants.py [--traces N] [--ticks T] [--attrs A] [--reps R]
Prints one line per variant: kernel time (CUDA events, best and mean of R after 2 warm-ups),
algorithmic GB/s and the fraction of the measured HBM copy peak. Inputs are >> L2, so no flush.
"""
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
    ap.add_argument("--traces", type=int, default=1 << 20)
    ap.add_argument("--ticks", type=int, default=2000)
    ap.add_argument("--attrs", type=int, default=1)
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--family", choices=["bands", "walk"], default="bands",
                    help="depth generator: SURVEY.md 8(d) family A (bands) or B (random walk, step 15, walls 0 and 500)")
    ap.add_argument("--full", type=str, default="all")
    ap.add_argument("--summary", type=str, default="all")
    args = ap.parse_args()
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    eng = ksa.ReplayEngine(0)
    N, T, A = args.traces, args.ticks, args.attrs
    if A == 1:
        pol = ksa.Policy(attr_mask=1)
        gen = ksa.GenParams(seed=0x5EED0001)
    else:
        pol = ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=250, scale_down_messages=25,
                         scale_up_pods=3, scale_down_pods=2, max_pods=50, min_pods=2,
                         scale_up_cool_down=12 * 10**9, scale_down_cool_down=45 * 10**9)
        gen = ksa.GenParams(seed=0x5EED0001, low_max=25, high_min=250, high_max=1000, overflow_log2=20)
    walk = None
    if args.family == "walk":
        gen = ksa.GenParams(seed=gen.seed, family=ksa.GEN_WALK, low_max=0, high_min=1, high_max=500, walk_step=15)
        walk = eng.gen_walk_init(gen, 0, N)
    depth = eng.gen_traces(gen, 0, T, A, 0, N, N, walk=walk)
    replicas = torch.empty((T, N), dtype=torch.int32, device=eng.dev)
    events = torch.empty((T, N), dtype=torch.uint8, device=eng.dev)
    state0 = eng.new_state(N, seed=gen.seed, lo=1, hi=5)
    summ = eng.new_summary(N)
    torch.cuda.synchronize()
    ref_hash = None

    def timed(fn):
        ts = []
        for it in range(2 + args.reps):
            st = state0.clone()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn(st)
            b.record()
            torch.cuda.synchronize()
            if it >= 2:
                ts.append(a.elapsed_time(b))
        return min(ts), sum(ts) / len(ts)

    def ids(spec, mode):
        if spec == "none":
            return []
        if spec != "all":
            return [int(x) for x in spec.split(",")]
        return [-2] + [i for i in range(1, 64) if eng.variant_name(mode, i)]

    print(f"# N={N} T={T} A={A} family={args.family} peak={peak} GB/s (measured copy)")
    bpd_full, bpd_sum = 8 * A + 5, 8 * A
    for vid in ids(args.full, 0):
        eng.set_variant(vid, 0)
        best, mean = timed(lambda st: eng.replay_full(pol, depth, T, A, N, N, 0, st, replicas=replicas, events=events))
        h = eng.traj_hash(replicas, events, T, N, N)
        if ref_hash is None:
            ref_hash = h.clone()
        same = bool(torch.equal(h, ref_hash))
        gbs = bpd_full * N * T / (best * 1e-3) / 1e9
        print(f"full    id={vid:3d} {eng.variant_name(0, 0 if vid == -2 else vid):24s} best={best:8.3f} ms mean={mean:8.3f} ms "
              f"{gbs:8.1f} GB/s frac={gbs / peak:5.3f} dec/s={N * T / (best * 1e-3):.3e} same_traj={same}", flush=True)
    for vid in ids(args.summary, 1):
        eng.set_variant(0, vid)

        def run(st):
            eng.lib.ksa_summary_clear(summ.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
            eng.replay_summary(pol, depth, T, A, N, N, 0, st, summary=summ)

        best, mean = timed(run)
        hs = summ.view(N, 64)[:, 56:64].contiguous().view(-1)
        same = ref_hash is None or bool(torch.equal(hs, ref_hash.view(torch.uint8)))
        gbs = bpd_sum * N * T / (best * 1e-3) / 1e9
        print(f"summary id={vid:3d} {eng.variant_name(1, 0 if vid == -2 else vid):24s} best={best:8.3f} ms mean={mean:8.3f} ms "
              f"{gbs:8.1f} GB/s frac={gbs / peak:5.3f} dec/s={N * T / (best * 1e-3):.3e} same_traj={same}", flush=True)
    eng.set_variant(0, 0)


if __name__ == "__main__":
    main()
