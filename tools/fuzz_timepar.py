#!/usr/bin/env python
"""Random shapes, policies and depths through the four replay entry points with the chunked paths
forced (csrc/ksa_timepar.cu), against the CPU oracle, bit for bit.

    python tools/fuzz_timepar.py [--cases 300] [--seed 1]

Depth families are mixed per trace: independent values, random walks, saturated / idle / quiet
stretches that start and end anywhere; policies are the adversarial ones of the parity tests plus
ordinary ones (small cool-downs, limits a few steps apart), so that the shift, closed-form, second
replay, probe and give-up paths all get exercised. Not part of the pytest suite (minutes of GPU)."""
from __future__ import annotations

import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import kube_sqs_autoscaler_b200 as ksa  # noqa: E402
from tests import oracle_lib  # noqa: E402
from tests.test_gpu_parity import random_policy, random_state  # noqa: E402

S = 10**9


def depths(rng, T, A, ld, up, down):
    up = int(np.clip(up, -2**40, 2**40))      # (adversarial policies have thresholds near the int64 limits)
    down = int(np.clip(down, -2**40, 2**40))
    d = rng.integers(0, max(2, 4 * max(up, 1)), size=(T, A, ld)).astype(np.int64) // A
    for n in range(ld):
        kind = rng.integers(0, 6)
        if kind == 0:
            continue
        if kind == 1:  # random walk
            w = np.clip(np.cumsum(rng.integers(-15, 16, size=T)) + rng.integers(0, 500), 0, 500)
            d[:, :, n] = (w // A)[:, None]
            continue
        for _ in range(rng.integers(1, 4)):  # stretches
            a = int(rng.integers(0, T))
            b = int(min(T, a + rng.integers(1, T)))
            what = rng.integers(0, 3)
            val = (up + int(rng.integers(0, 50))) if what == 0 else (min(down, up - 1) - int(rng.integers(0, 3)) if what == 1 else (up + down) // 2)
            d[a:b, :, n] = 0
            d[a:b, 0, n] = val
    return np.ascontiguousarray(d)


def run(cases: int, seed: int, eng=None) -> int:
    """Returns the number of mismatching cases (each is printed)."""
    rng = np.random.default_rng(seed)
    eng = eng or ksa.ReplayEngine(0)
    bad = 0
    for case in range(cases):
        A = int(rng.integers(1, 4))
        if rng.integers(0, 2):
            pol = random_policy(rng, A)
        else:
            up = int(rng.integers(20, 300))
            pol = ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=up, scale_down_messages=int(rng.integers(0, up)),
                             scale_up_cool_down=int(rng.integers(0, 9)) * 5 * S, scale_down_cool_down=int(rng.integers(0, 14)) * 5 * S,
                             max_pods=int(rng.integers(3, 12)), min_pods=int(rng.integers(0, 3)),
                             scale_up_pods=int(rng.integers(1, 4)), scale_down_pods=int(rng.integers(1, 3)))
        auto = rng.integers(0, 8) == 0  # now and then a call long enough for the library's own choice of chunks
        T = int(rng.integers(4096, 7000)) if auto else int(rng.integers(40, 1500))
        N = int(rng.choice([64, 520, 2048])) if auto else int(rng.choice([1, 8, 33, 64, 96, 130, 256, 515]))
        ld = N if rng.integers(0, 2) else (N + 15) // 16 * 16
        tick0 = int(rng.choice([0, 0, 7, 10**6]))
        d = depths(rng, T, A, ld, int(pol.scale_up_messages), int(pol.scale_down_messages))
        fail = None
        if rng.integers(0, 3) == 0:
            p = float(rng.choice([0.002, 0.05, 0.3]))
            fail = ((rng.random((T, ld)) < p) * 1 + (rng.random((T, ld)) < p) * 2 + (rng.random((T, ld)) < p) * 4).astype(np.uint8)
        st = random_state(rng, N, tick0)
        if rng.integers(0, 2):
            st = (rng.integers(0, 12, size=N).astype(np.int32), st[1], st[2])
        ref = oracle_lib.replay(pol.to_c(), d, T, A, N, ld, tick0, *st, want_summary=True, fail=fail)
        env = {"KSA_TIMEPAR_CHUNKS": str(int(rng.choice([2, 3, 4, 7, 12, 30]))), "KSA_TIMEPAR_WARM": str(int(rng.choice([0, 8, 24, 64, 128]))),
               "KSA_TIMEPAR_WIDE": str(int(rng.integers(0, 2))), "KSA_TIMEPAR_PROBE": str(int(rng.integers(0, 2)))}
        if auto:
            env["KSA_TIMEPAR_CHUNKS"] = ""
        for k, v in env.items():
            if v == "":
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        dd = torch.from_numpy(d).to(eng.dev)
        ff = None if fail is None else torch.from_numpy(fail).to(eng.dev)
        s1, s2 = eng.state_from_host(*st), eng.state_from_host(*st)
        r, e = eng.replay_full(pol, dd, T, A, N, ld, tick0, s1, fail=ff)
        summ = eng.replay_summary(pol, dd, T, A, N, ld, tick0, s2, fail=ff).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
        what = []
        if not (np.array_equal(r.cpu().numpy()[:, :N], ref.replicas[:, :N]) and np.array_equal(e.cpu().numpy()[:, :N], ref.events[:, :N])):
            what.append("trajectory")
        if summ.tobytes() != ref.summary.tobytes():
            what.append("summary: " + ",".join(f for f in ksa.SUMMARY_DTYPE.names if not np.array_equal(summ[f], ref.summary[f])))
        for name, s in (("full", s1), ("summary", s2)):
            if not (np.array_equal(s.replicas.cpu().numpy(), ref.state_replicas) and
                    np.array_equal(s.last_up.cpu().numpy(), ref.state_last_up) and np.array_equal(s.last_down.cpu().numpy(), ref.state_last_down)):
                what.append(f"state after the {name} replay")
        if what:
            bad += 1
            print(f"MISMATCH case {case} ({'; '.join(what)}): A={A} T={T} N={N} ld={ld} tick0={tick0} mask={fail is not None} {env} {pol}",
                  flush=True)
    for k in ("KSA_TIMEPAR_CHUNKS", "KSA_TIMEPAR_WARM", "KSA_TIMEPAR_WIDE", "KSA_TIMEPAR_PROBE"):
        os.environ.pop(k, None)
    return bad


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=300)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args()
    bad = run(args.cases, args.seed)
    print(f"{args.cases} cases, {bad} mismatches")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
