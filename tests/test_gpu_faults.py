"""GPU parity for the *_faults entry points: the three error paths of Run (main.go:44-47, 57-60,
71-74) replayed from a per-(tick, trace) failure mask. Through the C ABI,
bit-exact against the CPU oracle, which itself is checked against the literal wall-clock restatement
and the independent Python restatement in tests/test_oracle_kats.py.
"""
from __future__ import annotations

import os

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from tests import oracle_lib
from tests.helpers import S, load_golden, policy_from_vector, variants_present
from tests.test_gpu_parity import adversarial_depth, random_policy, random_state

pytestmark = pytest.mark.gpu

DERIVED = [v for v in load_golden("derived_vectors.json")["vectors"] if "fail" in v]


def run_gpu(engine, pol, depth_np, fail_np, T, A, N, ld, tick0, st):
    import torch

    depth = torch.from_numpy(np.ascontiguousarray(depth_np)).to(engine.dev)
    fail = torch.from_numpy(np.ascontiguousarray(fail_np)).to(engine.dev)
    s1 = engine.state_from_host(*st)
    replicas, events = engine.replay_full(pol, depth, T, A, N, ld, tick0, s1, fail=fail)
    s2 = engine.state_from_host(*st)
    summ = engine.replay_summary(pol, depth, T, A, N, ld, tick0, s2, fail=fail)
    h = engine.traj_hash(replicas, events, T, N, ld)
    torch.cuda.synchronize()
    return (replicas.cpu().numpy(), events.cpu().numpy(), s1, s2,
            summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1), h.cpu().numpy().view(np.uint64))


def check(out, ref, N, label):
    replicas, events, s1, s2, summ, h = out
    assert np.array_equal(replicas[:, :N], ref.replicas[:, :N]), f"replicas differ {label}"
    assert np.array_equal(events[:, :N], ref.events[:, :N]), f"events differ {label}"
    for s in (s1, s2):
        assert np.array_equal(s.replicas.cpu().numpy(), ref.state_replicas), label
        assert np.array_equal(s.last_up.cpu().numpy(), ref.state_last_up), label
        assert np.array_equal(s.last_down.cpu().numpy(), ref.state_last_down), label
    assert summ.tobytes() == ref.summary.tobytes(), f"summary differs {label}"
    assert np.array_equal(h, ref.summary["hash"]), label


@pytest.mark.parametrize("v", DERIVED, ids=lambda v: v["id"])
def test_q10_vectors_on_gpu(engine, v):
    pol = policy_from_vector(v)
    rows = np.asarray(v["depth"], dtype=np.int64)
    T, A = rows.shape
    for N, ld in ((1, 1), (19, 32)):
        depth = np.zeros((T, A, ld), dtype=np.int64)
        depth[:, :, :N] = rows[:, :, None]
        fail = np.zeros((T, ld), dtype=np.uint8)
        fail[:, :N] = np.asarray(v["fail"], dtype=np.uint8)[:, None]
        st = (np.full(N, v["r0"], np.int32), np.zeros(N, np.int64), np.zeros(N, np.int64))
        replicas, events, *_ = run_gpu(engine, pol, depth, fail, T, A, N, ld, 0, st)
        for n in range(N):
            assert replicas[:, n].tolist() == v["replicas"], v["id"]
            assert events[:, n].tolist() == v["events"], v["id"]


@pytest.mark.parametrize("A", [1, 2, 3])
def test_random_policies_with_failures(engine, A):
    rng = np.random.default_rng(900 + A)
    try:
        for case in range(8):
            pol = random_policy(rng, A)
            T = int(rng.integers(1, 80))
            N = int(rng.choice([1, 7, 64, 130, 1000, 2051]))
            ld = (N + 15) // 16 * 16 if case % 3 else N
            tick0 = int(rng.choice([0, 0, 23, 10**6, 2**40]))
            depth = adversarial_depth(rng, T, A, ld, pol)
            fail = rng.choice(np.array([0, 0, 0, 0, 1, 2, 4, 3, 5, 6, 7], dtype=np.uint8), size=(T, ld))
            st = random_state(rng, N, tick0)
            ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, tick0, *st, want_summary=True, fail=fail)
            for scalar in (False, True):
                engine.set_variant(-2 if scalar else 0, -2 if scalar else 0)
                out = run_gpu(engine, pol, depth, fail, T, A, N, ld, tick0, st)
                check(out, ref, N, f"A={A} case={case} N={N} T={T} ld={ld} scalar={scalar}")
            # every entry of the two faults variant tables the loaded build has (an id that only the
            # other mode's table has selects this mode's default)
            for fid in sorted(set(variants_present(engine, 2)) | set(variants_present(engine, 3))):
                fid = str(fid)
                os.environ["KSA_FAULTS_VARIANT"] = fid
                out = run_gpu(engine, pol, depth, fail, T, A, N, ld, tick0, st)
                check(out, ref, N, f"A={A} case={case} N={N} T={T} ld={ld} faults_variant={fid}")
            os.environ.pop("KSA_FAULTS_VARIANT", None)
    finally:
        os.environ.pop("KSA_FAULTS_VARIANT", None)
        engine.set_variant(0, 0)


def test_zero_mask_equals_the_plain_entry_points(engine):
    """With no failure bits set the *_faults kernels must reproduce ksa_replay_full / _summary."""
    import torch

    pol = ksa.Policy(attr_mask=1)
    gen = ksa.GenParams(seed=5)
    T, N = 300, 4096
    depth = engine.gen_traces(gen, 0, T, 1, 0, N, N)
    fail = torch.zeros((T, N), dtype=torch.uint8, device=engine.dev)
    s0 = engine.new_state(N, seed=5, lo=1, hi=5)
    a, b = s0.clone(), s0.clone()
    r1, e1 = engine.replay_full(pol, depth, T, 1, N, N, 0, a)
    r2, e2 = engine.replay_full(pol, depth, T, 1, N, N, 0, b, fail=fail)
    assert torch.equal(r1, r2) and torch.equal(e1, e2) and torch.equal(a.last_up, b.last_up)
    c, d = s0.clone(), s0.clone()
    s1 = engine.replay_summary(pol, depth, T, 1, N, N, 0, c)
    s2 = engine.replay_summary(pol, depth, T, 1, N, N, 0, d, fail=fail)
    assert torch.equal(s1, s2)


def test_generated_mask_chunks_and_counters(engine):
    """Config-2-shaped block with ~3 % failed polls and ~6 % failing scale calls: device generator ==
    host generator, T-chunked replay == one shot, and the failure counters add up."""
    pol = ksa.Policy(attr_mask=1)
    gen = ksa.GenParams(seed=0x5EED0001)
    T, N, ld = 600, 5000, 5008
    depth = engine.gen_traces(gen, 0, T, 1, 0, N, ld)
    fail = engine.gen_faults(gen.seed, 5, 4, 4, 0, T, 0, N, ld)
    f_host = oracle_lib.gen_faults(gen.seed, 5, 4, 4, 0, T, 0, N, ld)
    assert np.array_equal(fail.cpu().numpy()[:, :N], f_host[:, :N])
    assert 0.02 < (f_host[:, :N] & 1).mean() < 0.045
    st = oracle_lib.fresh_state(gen.seed, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth.cpu().numpy(), T, 1, N, ld, 0, *st, want_summary=True, fail=f_host)
    state = engine.new_state(N, seed=gen.seed, lo=1, hi=5)
    state_s = state.clone()
    summ = engine.new_summary(N)
    reps = []
    for t0 in range(0, T, 150):
        r, _ = engine.replay_full(pol, depth[t0:t0 + 150], 150, 1, N, ld, t0, state, fail=fail[t0:t0 + 150])
        engine.replay_summary(pol, depth[t0:t0 + 150], 150, 1, N, ld, t0, state_s, summary=summ,
                              fail=fail[t0:t0 + 150])
        reps.append(r.cpu().numpy())
    assert np.array_equal(np.concatenate(reps)[:, :N], ref.replicas[:, :N])
    got = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
    assert got.tobytes() == ref.summary.tobytes()
    assert got["poll_failed"].sum() == int((f_host[:, :N] & 1).sum())
    assert got["up_failed"].sum() > 0 and got["down_failed"].sum() > 0
    # every tick has exactly one poll; a failed poll is the only event of its tick
    ev = ref.events[:, :N]
    assert ((ev == 0xC0) == ((f_host[:, :N] & 1) == 1)).all()


@pytest.mark.parametrize("A,N,T,chunk", [(1, 1000, 130, 32), (3, 517, 77, 77), (2, 2049, 64, 7)])
def test_host_pipeline_with_failure_mask_matches_oracle(engine, A, N, T, chunk):
    """ksa_pipeline_replay_faults: host depths + host failure mask, chunked and overlapped, equal to
    the oracle's replay with the same mask (outputs, summary, carry state); two calls in a row
    (state carried through the host arrays) equal one call over all ticks."""
    rng = np.random.default_rng(A * 100 + N)
    pol = ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=250, scale_down_messages=25,
                     scale_up_cool_down=12 * S, scale_down_cool_down=45 * S, max_pods=9)
    depth = ksa.pinned_empty((T, A, N), np.int64)
    depth[...] = rng.integers(0, 200, size=(T, A, N), dtype=np.int64)
    fail = ksa.pinned_empty((T, N), np.uint8)
    fail[...] = rng.choice(np.array([0, 0, 0, 0, 1, 2, 4, 6, 7], dtype=np.uint8), size=(T, N))
    st = oracle_lib.fresh_state(8, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, N, 0, *st, want_summary=True, fail=fail)
    pipe = ksa.HostPipeline(0, N + 5, A, chunk)
    try:
        r, ju, jd = (np.array(x, copy=True) for x in st)
        replicas = ksa.pinned_empty((T, N), np.int32)
        events = ksa.pinned_empty((T, N), np.uint8)
        summary = np.empty(N, dtype=ksa.SUMMARY_DTYPE)
        pipe.replay(pol, depth, 0, r, ju, jd, replicas, events, summary, fail=fail)
        assert np.array_equal(replicas, ref.replicas) and np.array_equal(events, ref.events)
        assert summary.tobytes() == ref.summary.tobytes()
        assert np.array_equal(r, ref.state_replicas) and np.array_equal(ju, ref.state_last_up)
        assert np.array_equal(jd, ref.state_last_down)
        # split in time: [0, T1) then [T1, T), state carried by the caller
        T1 = T // 3
        r2, ju2, jd2 = (np.array(x, copy=True) for x in st)
        rr, ee = ksa.pinned_empty((T, N), np.int32), ksa.pinned_empty((T, N), np.uint8)
        pipe.replay(pol, depth[:T1], 0, r2, ju2, jd2, rr[:T1], ee[:T1], None, fail=fail[:T1])
        pipe.replay(pol, depth[T1:], T1, r2, ju2, jd2, rr[T1:], ee[T1:], None, fail=fail[T1:])
        assert np.array_equal(rr, ref.replicas) and np.array_equal(ee, ref.events)
        assert np.array_equal(r2, ref.state_replicas)
        # summary-only
        r3, ju3, jd3 = (np.array(x, copy=True) for x in st)
        pipe.replay(pol, depth, 0, r3, ju3, jd3, None, None, summary, fail=fail)
        assert summary.tobytes() == ref.summary.tobytes() and np.array_equal(r3, ref.state_replicas)
    finally:
        pipe.close()


@pytest.mark.parametrize("mode", ["summary", "full"])
def test_cli_with_injected_failures_matches_oracle(engine, mode):
    """ksa-replay --poll-fail-log2/--up-fail-log2/--down-fail-log2: totals (including the three
    failure counters) and the fingerprint digest equal the oracle's replay of the same generated
    depths and failure mask."""
    import re
    import subprocess

    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    N, T = 3000, 200
    r = subprocess.run([cli, "--poll-period=5s", "--scale-up-cool-down=12s", "--scale-down-cool-down=20s",
                        "--max-pods=9", "--traces", str(N), "--ticks", str(T), "--mode", mode, "--seed", "91",
                        "--poll-fail-log2", "4", "--up-fail-log2", "2", "--down-fail-log2", "3"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    m = re.search(r"scale up successful (\d+) \| scale down successful (\d+) \| at max (\d+) \| at min (\d+) \| "
                  r"up cool-down skips (\d+) \| down cool-down skips (\d+)", r.stdout)
    f = re.search(r"failed to get SQS messages (\d+) \| failed scaling up (\d+) \| failed scaling down (\d+)", r.stdout)
    assert m and f, r.stdout
    got = [int(x) for x in m.groups() + f.groups()]
    digest = int(re.search(r"fingerprints ([0-9a-f]{16})", r.stdout).group(1), 16)
    pol = ksa.Policy(attr_mask=1, max_pods=9, scale_up_cool_down=12 * S, scale_down_cool_down=20 * S)
    gen = ksa.GenParams(seed=91, low_max=10, high_min=100, high_max=400)
    ld = (N + 15) // 16 * 16
    d_host = oracle_lib.gen_traces(gen.to_c(), 0, T, 1, 0, N, ld)
    fail = oracle_lib.gen_faults(91, 4, 2, 3, 0, T, 0, N, ld)
    ref = oracle_lib.replay(pol.to_c(), d_host, T, 1, N, ld, 0, *oracle_lib.fresh_state(91, 0, N, 1, 9),
                            want_full=False, want_summary=True, fail=fail)
    want = [int(ref.summary[k].sum()) for k in ("up_applied", "down_applied", "up_noop", "down_noop", "up_cool",
                                                "down_cool", "poll_failed", "up_failed", "down_failed")]
    assert got == want and min(want[6:]) > 0
    assert digest == int(np.bitwise_xor.reduce(ref.summary["hash"]))
