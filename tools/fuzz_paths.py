#!/usr/bin/env python
"""Random parity campaign over the paths tools/fuzz_timepar.py does not reach: the plain kernels in
every compiled variant (with and without a failure mask, in T-chunks with the carry state handed
on), the policy sweep (ksa_replay_sweep, ksa_replay_sweep_static) with its tile kinds, the sweep
reduction, and the host pipeline. Everything against the C oracle / the Python restatement, bit for
bit. The time-parallel path is switched off here (KSA_TIMEPAR_CHUNKS=0): that is the other tool's.

    python tools/fuzz_paths.py --cases 300 --seed 1      # on a GPU box; exit code 1 on a mismatch

Test infrastructure: loads oracle/ (the checker), like tests/."""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import kube_sqs_autoscaler_b200 as ksa  # noqa: E402
from oracle import policy_py  # noqa: E402
from tests import oracle_lib  # noqa: E402
from tests.helpers import S, variants_present  # noqa: E402
from tests.test_gpu_parity import adversarial_depth, random_policy  # noqa: E402


def any_state(rng, N, tick0):
    """Replica counts anywhere in int32 (a third at the ends), stamps anywhere up to tick0."""
    r = rng.integers(-3, 13, size=N).astype(np.int64)
    far = rng.integers(0, 3, size=N) == 0
    r[far] = rng.choice(np.array([2**31 - 1, -(2**31), 2**31 - 3, -(2**31) + 2, 2**30, -(2**30), 2**26, -(2**26) - 1]),
                        size=int(far.sum()))
    wild = rng.integers(0, 6, size=N) == 0
    r[wild] = rng.integers(-(2**31), 2**31, size=int(wild.sum()))
    return r.astype(np.int32), rng.integers(0, tick0 + 1, size=N).astype(np.int64), rng.integers(0, tick0 + 1, size=N).astype(np.int64)


def everyday_policy(rng, A):
    up = int(rng.integers(20, 400))
    return ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=up, scale_down_messages=int(rng.integers(0, up)),
                      scale_up_cool_down=int(rng.integers(0, 9)) * 5 * S, scale_down_cool_down=int(rng.integers(0, 14)) * 5 * S,
                      max_pods=int(rng.integers(3, 12)), min_pods=int(rng.integers(0, 3)),
                      scale_up_pods=int(rng.integers(1, 4)), scale_down_pods=int(rng.integers(1, 3)))


def some_policy(rng, A):
    return random_policy(rng, A) if rng.integers(0, 2) else everyday_policy(rng, A)


def some_mask(rng, T, ld):
    p = float(rng.choice([0.002, 0.05, 0.3]))
    return ((rng.random((T, ld)) < p) * 1 + (rng.random((T, ld)) < p) * 2 + (rng.random((T, ld)) < p) * 4).astype(np.uint8)


def case_plain(rng, eng, fvs, svs, xvs):
    """One call or a run of T-chunks through a random variant; full and summary; maybe a mask."""
    A = int(rng.integers(1, 4))
    pol = some_policy(rng, A)
    T = int(rng.integers(1, 260))
    N = int(rng.choice([1, 5, 31, 64, 129, 515, 1000, 2051, 4100]))
    ld = N if rng.integers(0, 2) else (N + 15) // 16 * 16
    tick0 = int(rng.choice([0, 0, 17, 10**6, 2**40]))
    depth = adversarial_depth(rng, T, A, ld, pol)
    fail = some_mask(rng, T, ld) if rng.integers(0, 3) == 0 else None
    st = any_state(rng, N, tick0)
    ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, tick0, *st, want_summary=True, fail=fail)
    fv, sv = int(rng.choice(fvs)), int(rng.choice(svs))
    fault_variant = int(rng.choice(xvs))  # (0: the library's own choice; an id of the other mode's table means that too)
    if fail is not None and fault_variant:
        os.environ["KSA_FAULTS_VARIANT"] = str(fault_variant)
    eng.set_variant(fv, sv)
    try:
        cuts = sorted(set([0, T] + [int(x) for x in rng.integers(0, T + 1, size=int(rng.integers(0, 4)))]))
        d = torch.from_numpy(depth).to(eng.dev)
        f = None if fail is None else torch.from_numpy(fail).to(eng.dev)
        s1, s2 = eng.state_from_host(*st), eng.state_from_host(*st)
        summ = eng.new_summary(N)
        r = torch.empty((T, ld), dtype=torch.int32, device=eng.dev)
        e = torch.empty((T, ld), dtype=torch.uint8, device=eng.dev)
        for a, b in zip(cuts[:-1], cuts[1:]):
            if b == a:
                continue
            eng.replay_full(pol, d[a:b], b - a, A, N, ld, tick0 + a, s1, replicas=r[a:b], events=e[a:b],
                            fail=None if f is None else f[a:b])
            eng.replay_summary(pol, d[a:b], b - a, A, N, ld, tick0 + a, s2, summary=summ, fail=None if f is None else f[a:b])
    finally:
        eng.set_variant(0, 0)
        os.environ.pop("KSA_FAULTS_VARIANT", None)
    what = []
    if not np.array_equal(r.cpu().numpy()[:, :N], ref.replicas[:, :N]):
        what.append("replicas")
    if not np.array_equal(e.cpu().numpy()[:, :N], ref.events[:, :N]):
        what.append("events")
    got = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
    if got.tobytes() != ref.summary.tobytes():
        what.append("summary: " + ",".join(n for n in ksa.SUMMARY_DTYPE.names if not np.array_equal(got[n], ref.summary[n])))
    for name, s in (("full", s1), ("summary", s2)):
        if not (np.array_equal(s.replicas.cpu().numpy(), ref.state_replicas) and np.array_equal(s.last_up.cpu().numpy(), ref.state_last_up)
                and np.array_equal(s.last_down.cpu().numpy(), ref.state_last_down)):
            what.append(f"state after {name}")
    return what, f"plain A={A} T={T} N={N} ld={ld} tick0={tick0} cuts={cuts} mask={fail is not None} variants={fv}/{sv}/{fault_variant} {pol}"


def case_sweep(rng, eng):
    """A grid of policies laid out in runs that share thresholds (or not), replica counts anywhere."""
    A = int(rng.integers(1, 4))
    T = int(rng.integers(1, 140))
    N = int(rng.choice([1, 33, 130, 257, 600]))
    ld = N if rng.integers(0, 2) else (N + 15) // 16 * 16
    pols = []
    mask = some_policy(rng, A).attr_mask  # one grid, one set of attributes (ksa_replay_sweep's contract)
    want_p = int(rng.integers(1, 40))
    while len(pols) < want_p:
        base = some_policy(rng, A)
        for _ in range(int(rng.choice([1, 1, 3, 4, 8, 9]))):  # a run with the same thresholds and attributes
            q = some_policy(rng, A)
            pols.append(ksa.Policy(attr_mask=mask, scale_up_messages=base.scale_up_messages,
                                   scale_down_messages=base.scale_down_messages, poll_period=base.poll_period,
                                   scale_up_cool_down=q.scale_up_cool_down, scale_down_cool_down=q.scale_down_cool_down,
                                   scale_up_pods=q.scale_up_pods, scale_down_pods=q.scale_down_pods, max_pods=q.max_pods,
                                   min_pods=q.min_pods))
    P = len(pols)
    depth = adversarial_depth(rng, T, A, ld, pols[int(rng.integers(0, P))])
    r0 = any_state(rng, N, 0)[0]
    st = (r0, np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
    d = torch.from_numpy(depth).to(eng.dev)
    r0d = torch.from_numpy(r0).to(eng.dev)
    env = [{}, {}, {"KSA_SWEEP_NOSHARED": "1"}, {"KSA_SWEEP_GENERAL": "1"}][int(rng.integers(0, 4))]
    os.environ.update(env)
    try:
        if rng.integers(0, 3) == 0:
            got = eng.replay_sweep_static(pols, d, T, A, N, ld, r0d, torch.empty((P, N, 64), dtype=torch.uint8, device=eng.dev))
        else:
            got = eng.replay_sweep(pols, d, T, A, N, ld, r0d)
    finally:
        for k in env:
            os.environ.pop(k, None)
    w = [int(x) for x in rng.integers(-3, 9, size=4)]
    cost, best, best_cost = eng.sweep_reduce(got, P, N, *w)
    got = got.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
    what = []
    want = np.empty((P, N), dtype=ksa.SUMMARY_DTYPE)
    for i, pol in enumerate(pols):
        want[i] = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, 0, *st, want_full=False, want_summary=True).summary
        if got[i].tobytes() != want[i].tobytes():
            what.append(f"policy {i}: " + ",".join(n for n in ksa.SUMMARY_DTYPE.names if not np.array_equal(got[i][n], want[i][n])))
    rc, rb, rbc = policy_py.sweep_reduce(want, *w)
    if not (np.array_equal(cost.cpu().numpy(), rc) and np.array_equal(best.cpu().numpy(), rb) and np.array_equal(best_cost.cpu().numpy(), rbc)):
        what.append("reduction")
    return what[:4], f"sweep A={A} T={T} N={N} ld={ld} P={P} env={env} weights={w} first={pols[0]}"


def case_pipeline(rng, dev):
    A = int(rng.integers(1, 4))
    pol = some_policy(rng, A)
    T = int(rng.integers(1, 300))
    N = int(rng.choice([1, 33, 130, 1000, 2051]))
    chunk = int(rng.choice([1, 7, 16, 25, 64, 400]))
    tick0 = int(rng.choice([0, 17, 10**6]))
    depth = np.ascontiguousarray(adversarial_depth(rng, T, A, N, pol))
    fail = some_mask(rng, T, N) if rng.integers(0, 3) == 0 else None
    st = any_state(rng, N, tick0)
    ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, N, tick0, *st, want_summary=True, fail=fail)
    pipe = ksa.HostPipeline(dev, N, A, chunk)
    try:
        sr, su, sd = st[0].copy(), st[1].copy(), st[2].copy()
        r = np.empty((T, N), dtype=np.int32)
        e = np.empty((T, N), dtype=np.uint8)
        summ = np.zeros(N, dtype=ksa.SUMMARY_DTYPE)
        summ["min_replicas"] = 2**31 - 1
        summ["max_replicas"] = -(2**31)
        pipe.replay(pol, depth, tick0, sr, su, sd, replicas=r, events=e, summary=summ, fail=fail)
    finally:
        pipe.close()
    what = []
    if not np.array_equal(r, ref.replicas[:, :N]):
        what.append("replicas")
    if not np.array_equal(e, ref.events[:, :N]):
        what.append("events")
    if summ.tobytes() != ref.summary.tobytes():
        what.append("summary: " + ",".join(n for n in ksa.SUMMARY_DTYPE.names if not np.array_equal(summ[n], ref.summary[n])))
    if not (np.array_equal(sr, ref.state_replicas) and np.array_equal(su, ref.state_last_up) and np.array_equal(sd, ref.state_last_down)):
        what.append("state")
    return what, f"pipeline A={A} T={T} N={N} chunk={chunk} tick0={tick0} mask={fail is not None} {pol}"


def run(cases: int, seed: int, eng=None) -> int:
    """Returns the number of mismatching cases (each is printed)."""
    rng = np.random.default_rng(seed)
    eng = eng or ksa.ReplayEngine(0)
    fvs = [-2] + variants_present(eng, 0)
    svs = [-2] + variants_present(eng, 1)
    xvs = [0, 0] + sorted(set(variants_present(eng, 2)) | set(variants_present(eng, 3)))
    os.environ["KSA_TIMEPAR_CHUNKS"] = "0"
    bad = 0
    try:
        for case in range(cases):
            kind = int(rng.integers(0, 6))
            if kind < 3:
                what, label = case_plain(rng, eng, fvs, svs, xvs)
            elif kind < 5:
                what, label = case_sweep(rng, eng)
            else:
                what, label = case_pipeline(rng, eng.dev.index or 0)
            if what:
                bad += 1
                print(f"MISMATCH case {case} ({'; '.join(what)}): {label}", flush=True)
    finally:
        os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
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
