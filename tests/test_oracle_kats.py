"""Pin the oracle(s): the reference's own known-answer tests first, then the hand-derived quirk
vectors, then the two independent restatements against each other on random policies/traces.

CPU only. Nothing here touches libksareplay: this is the checker being checked.
"""
from __future__ import annotations

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from kube_sqs_autoscaler_b200 import Policy, render_log
from oracle import policy_py
from tests import oracle_lib
from tests.helpers import S, load_golden, policy_from_run_kat, policy_from_vector

KATS = load_golden("reference_kats.json")
DERIVED = load_golden("derived_vectors.json")


# ---- the reference's golden vectors (scale/scale_test.go, sqs/sqs_test.go, main_test.go) ---------

@pytest.mark.parametrize("k", KATS["scale"], ids=lambda k: k["name"])
def test_scale_kats_c_oracle(k):
    p = Policy(max_pods=k["max"], min_pods=k["min"], scale_up_pods=k["up_pods"],
               scale_down_pods=k["down_pods"]).to_c()
    assert oracle_lib.scale_calls(p, k["init"], k["calls"]) == k["expect"]


@pytest.mark.parametrize("k", KATS["scale"], ids=lambda k: k["name"])
def test_scale_kats_python_restatement(k):
    pas = policy_py.PodAutoScaler(k["max"], k["min"], k["up_pods"], k["down_pods"],
                                  policy_py.FakeDeployment(k["init"]))
    got = []
    for c in k["calls"]:
        (pas.scale_up if c == "up" else pas.scale_down)()
        got.append(pas.deployment.replicas)
    assert got == k["expect"]


@pytest.mark.parametrize("k", KATS["sum"], ids=lambda k: k["name"])
def test_sum_kat(k):
    assert oracle_lib.num_messages(k["attributes"]) == k["expect"]
    mock = policy_py.MockSQS({n: str(v) for n, v in zip(policy_py.DEFAULT_ATTRIBUTE_NAMES, k["attributes"])})
    assert policy_py.SqsClient(mock).num_messages() == k["expect"]


@pytest.mark.parametrize("k", KATS["run"], ids=lambda k: k["name"])
def test_run_kats_all_oracles(k):
    pol = policy_from_run_kat(k)
    for T in k["ticks"]:
        rows = np.tile(np.asarray(k["attributes"], dtype=np.int64), (T, 1))
        # literal wall-clock restatement (C)
        reps, _, final, _, _ = oracle_lib.run_wallclock(pol.to_c(), rows, k["init"])
        assert final == k["expect_final"] and reps[-1] == k["expect_final"], (k["name"], T, reps)
        # integer-tick batch restatement (C), N=1
        depth = rows.reshape(T, 3, 1)
        res = oracle_lib.replay(pol.to_c(), depth, T, 3, 1, 1, 0, [k["init"]], [0], [0])
        assert res.state_replicas[0] == k["expect_final"]
        assert np.array_equal(res.replicas[:, 0], reps)
        # independent Python restatement
        pr, _, _, _ = policy_py.replay_trace(pol.as_abi_dict(), [tuple(r) for r in rows.tolist()], k["init"])
        assert pr[-1] == k["expect_final"] and pr == reps.tolist()


# ---- derived quirk vectors (unpinned by the reference; hand-checked) -----------------------------

@pytest.mark.parametrize("v", DERIVED["vectors"], ids=lambda v: v["id"])
def test_derived_vectors_all_oracles(v):
    pol = policy_from_vector(v)
    rows = np.asarray(v["depth"], dtype=np.int64)
    T, A = rows.shape
    assert pol.num_attrs == A
    fail = np.asarray(v["fail"], dtype=np.uint8) if "fail" in v else None
    reps, evs, _, _, _ = oracle_lib.run_wallclock(pol.to_c(), rows, v["r0"], fail=fail)
    assert reps.tolist() == v["replicas"], "wall-clock C oracle"
    assert evs.tolist() == v["events"], "wall-clock C oracle"
    res = oracle_lib.replay(pol.to_c(), rows.reshape(T, A, 1), T, A, 1, 1, 0, [v["r0"]], [0], [0],
                            fail=None if fail is None else fail.reshape(T, 1))
    assert res.replicas[:, 0].tolist() == v["replicas"], "tick-model C oracle"
    assert res.events[:, 0].tolist() == v["events"], "tick-model C oracle"
    pr, pe, _, _ = policy_py.replay_trace(pol.as_abi_dict(), [tuple(r) for r in rows.tolist()], v["r0"],
                                          fail_rows=v.get("fail"))
    assert pr == v["replicas"] and pe == v["events"], "python restatement"


@pytest.mark.parametrize("c", DERIVED["cool_ticks"], ids=lambda c: f"{c['cool_ns']}/{c['poll_ns']}")
def test_cool_ticks(c):
    p = Policy(poll_period=c["poll_ns"], scale_up_cool_down=c["cool_ns"], scale_down_cool_down=c["cool_ns"])
    assert oracle_lib.cool_ticks(p.to_c()) == (c["ticks"], c["ticks"])


def test_cool_ticks_rejects_nonpositive_poll():
    with pytest.raises(ValueError):
        oracle_lib.cool_ticks(Policy(poll_period=0).to_c())


# ---- two restatements against each other ---------------------------------------------------------

I64 = st.integers(min_value=-(2**63), max_value=2**63 - 1)
SMALL = st.integers(min_value=-5, max_value=600)
near = st.one_of(SMALL, SMALL, SMALL, I64, st.sampled_from([2**63 - 1, -(2**63), 2**31, 2**32 + 3, -(2**31) - 1]))


@st.composite
def policies(draw):
    poll = draw(st.sampled_from([1, 3, S, 5 * S, 7 * S]))
    cool = st.one_of(st.integers(-2, 6).map(lambda k: k * poll), st.integers(0, 40 * S),
                     st.sampled_from([0, 2**63 - 1]))
    pods = st.one_of(st.integers(-3, 12), st.sampled_from([2**31 - 1, -(2**31), 2**32 + 2, 2**40 + 7]))
    return Policy(poll_period=poll, scale_up_cool_down=draw(cool), scale_down_cool_down=draw(cool),
                  scale_up_messages=draw(near), scale_down_messages=draw(near),
                  scale_up_pods=draw(pods), scale_down_pods=draw(pods), max_pods=draw(pods),
                  min_pods=draw(pods), attr_mask=draw(st.sampled_from([1, 2, 4, 3, 5, 6, 7])))


@settings(max_examples=300, deadline=None)
@given(pol=policies(), data=st.data())
def test_c_oracles_equal_python_restatement(pol, data):
    A = pol.num_attrs
    T = data.draw(st.integers(1, 24))
    rows = [tuple(data.draw(near) for _ in range(A)) for _ in range(T)]
    r0 = data.draw(st.one_of(st.integers(-3, 12), st.sampled_from([2**31 - 1, -(2**31), 2147483000])))
    tick0 = data.draw(st.sampled_from([0, 0, 5, 10**6, 2**40]))
    ju = data.draw(st.integers(0, tick0)) if tick0 else 0
    jd = data.draw(st.integers(0, tick0)) if tick0 else 0
    arr = np.asarray(rows, dtype=np.int64).reshape(T, A)
    fail_rows = None
    if data.draw(st.booleans()):  # half of the cases also inject call failures (Q10)
        fail_rows = [data.draw(st.sampled_from([0, 0, 0, 1, 2, 4, 6, 3, 7])) for _ in range(T)]
    fail = None if fail_rows is None else np.asarray(fail_rows, dtype=np.uint8)
    w_r, w_e, w_final, w_ju, w_jd = oracle_lib.run_wallclock(pol.to_c(), arr, r0, tick0, ju, jd, fail=fail)
    res = oracle_lib.replay(pol.to_c(), arr.reshape(T, A, 1), T, A, 1, 1, tick0, [r0], [ju], [jd],
                            want_summary=True, fail=None if fail is None else fail.reshape(T, 1))
    text = []
    p_r, p_e, p_ju, p_jd = policy_py.replay_trace(pol.as_abi_dict(), rows, r0, tick0, ju, jd,
                                                  fail_rows=fail_rows, text_log=text)
    assert w_r.tolist() == p_r and w_e.tolist() == p_e
    # the log lines the reference would print, reconstructed from the replay's outputs alone
    # (kube_sqs_autoscaler_b200.render_log), equal the ones the restated objects emitted
    sums = [((sum(r) + 2**63) % 2**64) - 2**63 for r in rows]
    assert render_log(pol, r0, sums, res.replicas[:, 0], res.events[:, 0]) == text
    assert res.replicas[:, 0].tolist() == p_r and res.events[:, 0].tolist() == p_e
    assert (w_ju, w_jd) == (p_ju, p_jd) == (int(res.state_last_up[0]), int(res.state_last_down[0]))
    assert w_final == p_r[-1]
    # summary consistency
    s = res.summary[0]
    assert s["final_replicas"] == p_r[-1]
    assert s["min_replicas"] == min(p_r) and s["max_replicas"] == max(p_r)
    for bit, name in ((1, "up_applied"), (2, "down_applied"), (4, "up_noop"), (8, "down_noop"),
                      (16, "up_cool"), (32, "down_cool")):
        assert s[name] == sum(1 for e in p_e if e & bit)
    assert s["poll_failed"] == sum(1 for e in p_e if e == 0xC0)
    assert s["up_failed"] == sum(1 for e in p_e if e != 0xC0 and e & 0x40)
    assert s["down_failed"] == sum(1 for e in p_e if e != 0xC0 and e & 0x80)
    lo = hi = 0  # the two lanes of ksa_summary.hash (include/ksa_replay.h)
    for r, e in zip(p_r, p_e):
        wide = sum(((e >> k) & 1) << pos for k, pos in enumerate((0, 5, 10, 15, 20, 25, 30, 31)))
        lo = (lo * 0x9E3779B1 + (r & 0xFFFFFFFF)) & 0xFFFFFFFF
        hi = (hi * 0x85EBCA6B + wide) & 0xFFFFFFFF
    assert int(s["hash"]) == (hi << 32 | lo)
    assert int(s["sum_replicas"]) == ((sum(p_r) + 2**63) % 2**64) - 2**63


def test_chunked_replay_equals_one_shot():
    """Carry state makes T-chunking exact (needed for 1M x 10k x 3 attributes, which does not fit)."""
    rng = np.random.default_rng(7)
    pol = Policy(scale_up_messages=250, scale_down_messages=25, scale_up_pods=3, scale_down_pods=2,
                 max_pods=50, min_pods=2, scale_up_cool_down=12 * S, scale_down_cool_down=45 * S)
    T, A, N = 200, 3, 257
    depth = rng.integers(0, 140, size=(T, A, N), dtype=np.int64)
    st0 = oracle_lib.fresh_state(1, 0, N, 1, 5)
    one = oracle_lib.replay(pol.to_c(), depth, T, A, N, N, 0, *st0, want_summary=True)
    state, summ, reps = st0, None, []
    for t0 in range(0, T, 37):
        tc = min(37, T - t0)
        part = oracle_lib.replay(pol.to_c(), depth[t0:t0 + tc], tc, A, N, N, t0, *state,
                                 want_summary=True, summary=summ)
        state = (part.state_replicas, part.state_last_up, part.state_last_down)
        summ = part.summary
        reps.append(part.replicas)
    assert np.array_equal(np.concatenate(reps), one.replicas)
    assert summ.tobytes() == one.summary.tobytes()
    assert np.array_equal(state[1], one.state_last_up) and np.array_equal(state[2], one.state_last_down)


def test_oracle_threads_do_not_change_results():
    rng = np.random.default_rng(3)
    pol = Policy(attr_mask=1)
    T, N = 64, 5000
    depth = rng.integers(0, 400, size=(T, 1, N), dtype=np.int64)
    st0 = oracle_lib.fresh_state(9, 0, N, 1, 5)
    a = oracle_lib.replay(pol.to_c(), depth, T, 1, N, N, 0, *st0, want_summary=True, threads=1)
    b = oracle_lib.replay(pol.to_c(), depth, T, 1, N, N, 0, *st0, want_summary=True, threads=3)
    assert a.threads == 1 and b.threads == 3
    assert np.array_equal(a.replicas, b.replicas) and np.array_equal(a.events, b.events)
    assert a.summary.tobytes() == b.summary.tobytes()
    assert np.array_equal(oracle_lib.traj_hash(a.replicas, a.events, T, N, N), a.summary["hash"])


def test_wide_event_formula_matches_its_definition():
    """KSA_EV_WIDE (include/ksa_replay.h; also ksa_wide_event in csrc/ksa_policy.cuh): the
    multiply-and-mask spelling equals 'bit k to bit 5k, failure bits to 30 and 31' for every byte."""
    for ev in range(256):
        want = sum(((ev >> k) & 1) << pos for k, pos in enumerate((0, 5, 10, 15, 20, 25, 30, 31)))
        got = (((ev & 7) * 0x111) & 0x421) | ((((ev >> 3 & 7) * 0x111) & 0x421) << 15) | ((ev >> 6) << 30)
        assert got == want, ev
