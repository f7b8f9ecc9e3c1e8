"""GPU parity: the CUDA path, called through the C ABI (ctypes -> libksareplay.so), against the CPU
oracle on the same inputs. Integer work: the bar is bit-exact, everywhere.

Run with `pytest -m gpu` on the B200 box. Nothing here reads /root/reference.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from kube_sqs_autoscaler_b200 import _abi
from tests import oracle_lib
from tests.helpers import S, load_golden, policy_from_run_kat, policy_from_vector, variants_present

pytestmark = pytest.mark.gpu

KATS = load_golden("reference_kats.json")
DERIVED = load_golden("derived_vectors.json")


def full_variants(engine):
    """-2 = the scalar kernel, then every full-mode variant the loaded build has."""
    return [-2] + variants_present(engine, 0)


def summary_variants(engine):
    return [-2] + variants_present(engine, 1)  # 21.. ordered-thresholds step; 30.. bulk-copy staging (tuning build)


def torch_mod():
    import torch

    return torch


def gpu_full(engine, pol, depth_np, T, A, N, ld, tick0, st, want_summary=True):
    """Run full + summary replay on the GPU for host depth [T][A][ld]; returns numpy results."""
    torch = torch_mod()
    depth = torch.from_numpy(np.ascontiguousarray(depth_np)).to(engine.dev)
    state = engine.state_from_host(*st)
    replicas, events = engine.replay_full(pol, depth, T, A, N, ld, tick0, state)
    out = {
        "replicas": replicas.cpu().numpy(), "events": events.cpu().numpy(),
        "st_r": state.replicas.cpu().numpy(), "st_ju": state.last_up.cpu().numpy(),
        "st_jd": state.last_down.cpu().numpy(),
    }
    if want_summary:
        state2 = engine.state_from_host(*st)
        summ = engine.replay_summary(pol, depth, T, A, N, ld, tick0, state2)
        out["summary"] = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
        out["st2_r"] = state2.replicas.cpu().numpy()
        out["st2_ju"] = state2.last_up.cpu().numpy()
        out["st2_jd"] = state2.last_down.cpu().numpy()
        out["hash"] = engine.traj_hash(replicas, events, T, N, ld).cpu().numpy().view(np.uint64)
    return out


def assert_matches_oracle(out, ref, N, label=""):
    assert np.array_equal(out["replicas"][:, :N], ref.replicas[:, :N]), f"replicas differ {label}"
    assert np.array_equal(out["events"][:, :N], ref.events[:, :N]), f"events differ {label}"
    assert np.array_equal(out["st_r"], ref.state_replicas), f"state replicas differ {label}"
    assert np.array_equal(out["st_ju"], ref.state_last_up), f"last_up differs {label}"
    assert np.array_equal(out["st_jd"], ref.state_last_down), f"last_down differs {label}"
    if "summary" in out:
        assert out["summary"].tobytes() == ref.summary.tobytes(), f"summary differs {label}"
        assert np.array_equal(out["st2_r"], ref.state_replicas)
        assert np.array_equal(out["st2_ju"], ref.state_last_up)
        assert np.array_equal(out["st2_jd"], ref.state_last_down)
        assert np.array_equal(out["hash"], ref.summary["hash"]), f"traj_hash differs {label}"


# ---- config 1: the reference's own tests, through the kernels ------------------------------------

@pytest.mark.parametrize("k", KATS["run"], ids=lambda k: k["name"])
def test_reference_run_kats_on_gpu(engine, k):
    pol = policy_from_run_kat(k)
    for T in k["ticks"]:
        for N, ld in ((1, 1), (1, 16), (37, 48)):  # N=1 as in the reference, and padded / replicated
            depth = np.zeros((T, 3, ld), dtype=np.int64)
            depth[:, :, :N] = np.asarray(k["attributes"], dtype=np.int64)[None, :, None]
            st = (np.full(N, k["init"], np.int32), np.zeros(N, np.int64), np.zeros(N, np.int64))
            out = gpu_full(engine, pol, depth, T, 3, N, ld, 0, st)
            assert (out["st_r"] == k["expect_final"]).all(), (k["name"], T, N)
            assert (out["replicas"][-1, :N] == k["expect_final"]).all()
            assert (out["summary"]["final_replicas"] == k["expect_final"]).all()


@pytest.mark.parametrize("k", KATS["scale"], ids=lambda k: k["name"])
def test_reference_scale_kats_on_gpu(engine, k):
    """ScaleUp/ScaleDown called directly (scale/scale_test.go) == no cool-down, depth forced."""
    pol = ksa.Policy(max_pods=k["max"], min_pods=k["min"], scale_up_pods=k["up_pods"],
                     scale_down_pods=k["down_pods"], scale_up_cool_down=0, scale_down_cool_down=0,
                     scale_up_messages=100, scale_down_messages=10, attr_mask=1)
    rows = np.asarray([[1000 if c == "up" else 0] for c in k["calls"]], dtype=np.int64)
    T = len(k["calls"])
    st = (np.array([k["init"]], np.int32), np.zeros(1, np.int64), np.zeros(1, np.int64))
    out = gpu_full(engine, pol, rows.reshape(T, 1, 1), T, 1, 1, 1, 0, st)
    assert out["replicas"][:, 0].tolist() == k["expect"]


def test_reference_sum_kat_on_gpu(engine):
    """sqs/sqs_test.go:19-25: 10+10+10 = 30. Observable through the thresholds: 30 >= 30 scales up,
    30 < 31 does not."""
    k = KATS["sum"][0]
    rows = np.asarray([k["attributes"]], dtype=np.int64).reshape(1, 3, 1)
    st = (np.array([3], np.int32), np.zeros(1, np.int64), np.zeros(1, np.int64))
    for thr, expect in ((k["expect"], 4), (k["expect"] + 1, 3)):
        pol = ksa.Policy(scale_up_messages=thr, scale_down_messages=-1, scale_up_cool_down=0)
        out = gpu_full(engine, pol, rows, 1, 3, 1, 1, 0, st)
        assert out["replicas"][0, 0] == expect


@pytest.mark.parametrize("v", [v for v in DERIVED["vectors"] if "fail" not in v], ids=lambda v: v["id"])
def test_derived_quirk_vectors_on_gpu(engine, v):  # the Q10 (failure) vectors: tests/test_gpu_faults.py
    pol = policy_from_vector(v)
    rows = np.asarray(v["depth"], dtype=np.int64)
    T, A = rows.shape
    for N, ld in ((1, 1), (21, 32)):
        depth = np.zeros((T, A, ld), dtype=np.int64)
        depth[:, :, :N] = rows[:, :, None]
        st = (np.full(N, v["r0"], np.int32), np.zeros(N, np.int64), np.zeros(N, np.int64))
        out = gpu_full(engine, pol, depth, T, A, N, ld, 0, st)
        for n in range(N):
            assert out["replicas"][:, n].tolist() == v["replicas"], v["id"]
            assert out["events"][:, n].tolist() == v["events"], v["id"]


# ---- random policies x adversarial traces, every kernel variant ----------------------------------

def random_policy(rng, A):
    poll = int(rng.choice([1, 3, S, 5 * S, 7 * S]))
    def cool():
        c = rng.integers(0, 4)
        if c == 0:
            return int(rng.integers(-2, 7)) * poll
        if c == 1:
            return int(rng.integers(0, 40 * S))
        if c == 2:
            return 0
        return int(min(2**63 - 1, int(rng.choice([2**62, 2**31, 5])) * poll))
    def pods():
        if rng.integers(0, 5) == 0:
            return int(rng.choice([2**31 - 1, -(2**31), 2**32 + 2, 2**40 + 7, 0, -1]))
        return int(rng.integers(-3, 13))
    def thr():
        c = rng.integers(0, 6)
        if c < 4:
            return int(rng.integers(-5, 600))
        return int(rng.choice([2**63 - 1, -(2**63), 2**31, 2**32 + 3, -(2**31) - 1, 0]))
    mask = [m for m in (1, 2, 4, 3, 5, 6, 7) if bin(m).count("1") == A]
    return ksa.Policy(poll_period=poll, scale_up_cool_down=cool(), scale_down_cool_down=cool(),
                      scale_up_messages=thr(), scale_down_messages=thr(), scale_up_pods=pods(),
                      scale_down_pods=pods(), max_pods=pods(), min_pods=pods(),
                      attr_mask=int(rng.choice(mask)))


def adversarial_depth(rng, T, A, ld, pol):
    base = rng.integers(-5, 600, size=(T, A, ld), dtype=np.int64)
    special = np.array([2**63 - 1, -(2**63), 2**31, 2**32 + 3, -(2**31) - 1, 0,
                        pol.scale_up_messages, pol.scale_down_messages,
                        min(pol.scale_up_messages, 2**63 - 2) + 1, max(pol.scale_down_messages, -(2**63) + 1) - 1],
                       dtype=np.int64)
    pick = rng.integers(0, 5, size=base.shape) == 0
    base[pick] = special[rng.integers(0, special.size, size=int(pick.sum()))]
    return base


def random_state(rng, N, tick0):
    r = rng.integers(-3, 13, size=N).astype(np.int32)
    edge = rng.integers(0, 8, size=N) == 0
    r[edge] = rng.choice(np.array([2**31 - 1, -(2**31), 2147483000, -2147483000], dtype=np.int64),
                         size=int(edge.sum())).astype(np.int32)
    ju = rng.integers(0, tick0 + 1, size=N).astype(np.int64)
    jd = rng.integers(0, tick0 + 1, size=N).astype(np.int64)
    return r, ju, jd


@pytest.mark.parametrize("A", [1, 2, 3])
def test_random_policies_all_variants(engine, A):
    rng = np.random.default_rng(1234 + A)
    cases = []
    for case in range(6):
        pol = random_policy(rng, A)
        T = int(rng.integers(1, 70))
        N = int(rng.choice([1, 5, 64, 129, 1000, 2051]))
        ld = (N + 15) // 16 * 16 if case % 3 else N
        tick0 = int(rng.choice([0, 0, 17, 10**6, 2**40]))
        depth = adversarial_depth(rng, T, A, ld, pol)
        st = random_state(rng, N, tick0)
        ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, tick0, *st, want_summary=True)
        cases.append((pol, T, N, ld, tick0, depth, st, ref))
    try:
        fvs, svs = full_variants(engine), summary_variants(engine)
        for k in range(max(len(fvs), len(svs))):
            fv, sv = fvs[k % len(fvs)], svs[k % len(svs)]
            engine.set_variant(fv, sv)
            for pol, T, N, ld, tick0, depth, st, ref in cases:
                out = gpu_full(engine, pol, depth, T, A, N, ld, tick0, st)
                assert_matches_oracle(out, ref, N, f"A={A} full_variant={fv} summary_variant={sv} N={N} T={T} ld={ld}")
    finally:
        engine.set_variant(0, 0)


def test_default_policy_bands_large_block(engine):
    """Config-2-shaped block (default policy, single attribute, band generator) at a size the
    oracle finishes in seconds; every trace compared in full."""
    pol = ksa.Policy(attr_mask=1)
    gen = ksa.GenParams(seed=0x5EED0001)
    T, N = 512, 1 << 15
    depth = engine.gen_traces(gen, 0, T, 1, 0, N, N)
    state = engine.new_state(N, seed=gen.seed, lo=1, hi=5)
    replicas, events = engine.replay_full(pol, depth, T, 1, N, N, 0, state)
    d_host = depth.cpu().numpy()
    assert np.array_equal(d_host, oracle_lib.gen_traces(gen.to_c(), 0, T, 1, 0, N, N))
    ref = oracle_lib.replay(pol.to_c(), d_host, T, 1, N, N, 0,
                            *oracle_lib.fresh_state(gen.seed, 0, N, 1, 5), want_summary=True)
    assert np.array_equal(replicas.cpu().numpy(), ref.replicas)
    assert np.array_equal(events.cpu().numpy(), ref.events)
    assert np.array_equal(state.replicas.cpu().numpy(), ref.state_replicas)
    # the workload is not degenerate: every outcome occurs
    tot = {k: int(ref.summary[k].sum()) for k in ("up_applied", "down_applied", "up_noop", "down_noop", "up_cool", "down_cool")}
    assert all(v > 0 for v in tot.values()), tot


# ---- generator ------------------------------------------------------------------------------------

@pytest.mark.parametrize("family", [ksa.GEN_BANDS, ksa.GEN_WALK, ksa.GEN_ADVERSARIAL])
@pytest.mark.parametrize("A", [1, 2, 3])
def test_generator_matches_host_and_is_shard_chunk_invariant(engine, family, A):
    torch = torch_mod()
    gen = ksa.GenParams(seed=42 + family, family=family, low_max=25, high_min=250, high_max=1000,
                        walk_step=15, overflow_log2=6 if A == 3 else 0)
    T, N, ld = 97, 777, 784
    walk = engine.gen_walk_init(gen, 0, N) if family == ksa.GEN_WALK else None
    hwalk = oracle_lib.gen_walk_init(gen.to_c(), 0, N) if family == ksa.GEN_WALK else None
    if walk is not None:
        assert np.array_equal(walk.cpu().numpy(), hwalk)
    dev = engine.gen_traces(gen, 0, T, A, 0, N, ld, walk=walk).cpu().numpy()
    host = oracle_lib.gen_traces(gen.to_c(), 0, T, A, 0, N, ld, walk=hwalk)
    assert np.array_equal(dev[:, :, :N], host[:, :, :N])
    if walk is not None:
        assert np.array_equal(walk.cpu().numpy(), hwalk)
    if A == 3 and family == ksa.GEN_BANDS:
        assert (dev[:, 0, :N] == 2**63 - 1).any(), "overflow triples expected at rate 2^-6"
    # shard + chunk: traces [300, 500), ticks 41..97 generated on their own == the same block
    n0, n1, t0 = 300, 500, 40
    w2 = None
    if family == ksa.GEN_WALK:
        w2 = engine.gen_walk_init(gen, n0, n1 - n0)
        engine.gen_traces(gen, 0, t0, A, n0, n1 - n0, n1 - n0, walk=w2)  # advance the walk to tick t0
    part = engine.gen_traces(gen, t0, T - t0, A, n0, n1 - n0, n1 - n0, walk=w2).cpu().numpy()
    assert np.array_equal(part, dev[t0:, :, n0:n1])
    torch.cuda.synchronize()


def test_state_init_matches_host(engine):
    st = engine.new_state(1000, seed=0x5EED0001, lo=1, hi=5, trace0=12345)
    r, ju, jd = oracle_lib.fresh_state(0x5EED0001, 12345, 1000, 1, 5)
    assert np.array_equal(st.replicas.cpu().numpy(), r)
    assert (st.last_up.cpu().numpy() == 0).all() and (st.last_down.cpu().numpy() == 0).all()
    assert r.min() == 1 and r.max() == 5


# ---- layout edge cases ----------------------------------------------------------------------------

def test_unaligned_pointers_take_the_scalar_path(engine):
    """Base pointers that are only 8-byte aligned and an odd pitch: same results."""
    torch = torch_mod()
    rng = np.random.default_rng(5)
    pol = ksa.Policy(attr_mask=1)
    T, N, ld = 33, 301, 303
    depth_np = rng.integers(0, 400, size=(T, 1, ld), dtype=np.int64)
    st = oracle_lib.fresh_state(3, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth_np, T, 1, N, ld, 0, *st)
    big = torch.zeros(T * ld + 1, dtype=torch.int64, device=engine.dev)
    depth = big[1:]  # 8-byte aligned only
    depth.copy_(torch.from_numpy(depth_np.reshape(-1)).to(engine.dev))
    rbig = torch.zeros(T * ld + 1, dtype=torch.int32, device=engine.dev)
    ebig = torch.zeros(T * ld + 3, dtype=torch.uint8, device=engine.dev)
    state = engine.state_from_host(*st)
    pc, sc = pol.to_c(), state.to_c()
    _abi.check(engine.lib.ksa_replay_full(C.byref(pc), depth.data_ptr(), T, 1, N, ld, 0, C.byref(sc),
                                          rbig[1:].data_ptr(), ebig[3:].data_ptr(), None))
    torch.cuda.synchronize()
    assert np.array_equal(rbig[1:].cpu().numpy().reshape(T, ld)[:, :N], ref.replicas[:, :N])
    assert np.array_equal(ebig[3:].cpu().numpy().reshape(T, ld)[:, :N], ref.events[:, :N])


def test_padding_columns_are_never_written(engine):
    torch = torch_mod()
    pol = ksa.Policy(attr_mask=1)
    T, N, ld = 20, 100, 112
    depth = torch.full((T, 1, ld), 500, dtype=torch.int64, device=engine.dev)
    replicas = torch.full((T, ld), -77, dtype=torch.int32, device=engine.dev)
    events = torch.full((T, ld), 0xEE, dtype=torch.uint8, device=engine.dev)
    state = engine.new_state(N, seed=1, lo=1, hi=5)
    engine.replay_full(pol, depth, T, 1, N, ld, 0, state, replicas=replicas, events=events)
    assert (replicas[:, N:] == -77).all() and (events[:, N:] == 0xEE).all()
    assert (replicas[:, :N] != -77).all()


def test_optional_outputs(engine):
    torch = torch_mod()
    rng = np.random.default_rng(11)
    pol = ksa.Policy(attr_mask=1)
    T, N = 40, 515
    ld = 528
    depth_np = rng.integers(0, 400, size=(T, 1, ld), dtype=np.int64)
    st = oracle_lib.fresh_state(3, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth_np, T, 1, N, ld, 0, *st)
    depth = torch.from_numpy(depth_np).to(engine.dev)
    s1 = engine.state_from_host(*st)
    r_only, e_none = engine.replay_full(pol, depth, T, 1, N, ld, 0, s1, want_events=False)
    assert e_none is None and np.array_equal(r_only.cpu().numpy()[:, :N], ref.replicas[:, :N])
    s2 = engine.state_from_host(*st)
    r_none, e_only = engine.replay_full(pol, depth, T, 1, N, ld, 0, s2, want_replicas=False)
    assert r_none is None and np.array_equal(e_only.cpu().numpy()[:, :N], ref.events[:, :N])
    s3 = engine.state_from_host(*st)
    engine.replay_full(pol, depth, T, 1, N, ld, 0, s3, want_replicas=False, want_events=False)
    for s in (s1, s2, s3):
        assert np.array_equal(s.replicas.cpu().numpy(), ref.state_replicas)


# ---- chunking / carry state -----------------------------------------------------------------------

def test_t_chunks_with_carry_state_equal_one_shot(engine):
    """Config 3's execution shape: A=3, non-default policy, T-chunks carrying (r, ju, jd)."""
    pol = ksa.Policy(scale_up_messages=250, scale_down_messages=25, scale_up_pods=3, scale_down_pods=2,
                     max_pods=50, min_pods=2, scale_up_cool_down=12 * S, scale_down_cool_down=45 * S)
    gen = ksa.GenParams(seed=0x5EED0003, low_max=25, high_min=250, high_max=1000, overflow_log2=8)
    T, A, N, ld = 1000, 3, 4099, 4112
    depth = engine.gen_traces(gen, 0, T, A, 0, N, ld)
    d_host = depth.cpu().numpy()
    st = oracle_lib.fresh_state(gen.seed, 0, N, 2, 50)
    ref = oracle_lib.replay(pol.to_c(), d_host, T, A, N, ld, 0, *st, want_summary=True)
    state = engine.new_state(N, seed=gen.seed, lo=2, hi=50)
    state_s = state.clone()
    summ = engine.new_summary(N)
    reps, evs = [], []
    for t0 in range(0, T, 250):
        r, e = engine.replay_full(pol, depth[t0:t0 + 250], 250, A, N, ld, t0, state)
        engine.replay_summary(pol, depth[t0:t0 + 250], 250, A, N, ld, t0, state_s, summary=summ)
        reps.append(r.cpu().numpy())
        evs.append(e.cpu().numpy())
    assert np.array_equal(np.concatenate(reps)[:, :N], ref.replicas[:, :N])
    assert np.array_equal(np.concatenate(evs)[:, :N], ref.events[:, :N])
    assert summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.summary.tobytes()
    assert np.array_equal(state.last_up.cpu().numpy(), ref.state_last_up)
    assert np.array_equal(state_s.last_down.cpu().numpy(), ref.state_last_down)


def test_policy_change_between_chunks(engine):
    """Stamps are absolute ticks, so the policy may change at a chunk boundary (e.g. replaying a
    cool-down retune): chunk 2 under the new cool-down still honours stamps taken in chunk 1."""
    rng = np.random.default_rng(21)
    N, ld, T1, T2 = 300, 304, 25, 40
    p1 = ksa.Policy(attr_mask=1, scale_up_cool_down=10 * S, scale_down_cool_down=30 * S)
    p2 = ksa.Policy(attr_mask=1, scale_up_cool_down=60 * S, scale_down_cool_down=5 * S, max_pods=9)
    d = rng.integers(0, 400, size=(T1 + T2, 1, ld), dtype=np.int64)
    st = oracle_lib.fresh_state(5, 0, N, 1, 5)
    a = oracle_lib.replay(p1.to_c(), d[:T1], T1, 1, N, ld, 0, *st)
    b = oracle_lib.replay(p2.to_c(), d[T1:], T2, 1, N, ld, T1, a.state_replicas, a.state_last_up, a.state_last_down)
    torch = torch_mod()
    depth = torch.from_numpy(d).to(engine.dev)
    state = engine.state_from_host(*st)
    engine.replay_full(p1, depth[:T1], T1, 1, N, ld, 0, state)
    r2, e2 = engine.replay_full(p2, depth[T1:], T2, 1, N, ld, T1, state)
    assert np.array_equal(r2.cpu().numpy()[:, :N], b.replicas[:, :N])
    assert np.array_equal(e2.cpu().numpy()[:, :N], b.events[:, :N])
    assert np.array_equal(state.last_up.cpu().numpy(), b.state_last_up)


def test_cool_down_longer_than_any_chunk(engine):
    """cool = 2^62 ticks of 1 ns: the first scale-up never happens, stamps stay 0; and a trace
    that did scale at tick0 - 1 stays blocked. Exercises the KSA_COOL_CAP clamp."""
    torch = torch_mod()
    pol = ksa.Policy(attr_mask=1, poll_period=1, scale_up_cool_down=2**62, scale_down_cool_down=2**62)
    T, N = 50, 64
    depth = torch.full((T, 1, N), 500, dtype=torch.int64, device=engine.dev)
    depth[::2] = 0
    tick0 = 2**45
    st = (np.full(N, 3, np.int32), np.full(N, tick0 - 1, np.int64), np.zeros(N, np.int64))
    ref = oracle_lib.replay(pol.to_c(), depth.cpu().numpy(), T, 1, N, N, tick0, *st, want_summary=True)
    out = gpu_full(engine, pol, depth.cpu().numpy(), T, 1, N, N, tick0, st)
    assert_matches_oracle(out, ref, N)
    assert (out["replicas"] == 3).all() and (out["st_ju"] == tick0 - 1).all()


# ---- sweep ----------------------------------------------------------------------------------------

def test_policy_sweep_matches_oracle_per_policy(engine):
    """Config 4's shape at small scale: a grid including overlapping thresholds and min>max."""
    torch = torch_mod()
    gen = ksa.GenParams(seed=77, family=ksa.GEN_WALK, high_max=500, walk_step=15)
    T, N = 300, 130
    walk = engine.gen_walk_init(gen, 0, N)
    depth = engine.gen_traces(gen, 0, T, 1, 0, N, N, walk=walk)
    d_host = depth.cpu().numpy()
    pols = []
    for up in (100, 250, 30):
        for down in (10, 120):
            for cu, cd in ((10 * S, 30 * S), (0, 0), (12 * S, 45 * S)):
                for mn, mx, su, sd in ((1, 5, 1, 1), (2, 50, 3, 2), (7, 2, 1, 1)):
                    pols.append(ksa.Policy(attr_mask=1, scale_up_messages=up, scale_down_messages=down,
                                           scale_up_cool_down=cu, scale_down_cool_down=cd, min_pods=mn,
                                           max_pods=mx, scale_up_pods=su, scale_down_pods=sd))
    pols = pols[:53]  # not a multiple of the policy tile
    r0 = oracle_lib.fresh_state(gen.seed, 0, N, 1, 5)
    got = engine.replay_sweep(pols, depth, T, 1, N, N, torch.from_numpy(r0[0]).to(engine.dev))
    got = got.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(len(pols), N)
    for i, pol in enumerate(pols):
        ref = oracle_lib.replay(pol.to_c(), d_host, T, 1, N, N, 0, *r0, want_full=False, want_summary=True)
        assert got[i].tobytes() == ref.summary.tobytes(), f"policy {i}: {pol}"


def test_sweep_fast_path_edges_match_oracle(engine, monkeypatch):
    """The sweep's narrow kernels (all policies ordered, thresholds inside int32) compare the
    attribute sum saturated to int32. Grid with thresholds at the edges of the allowed range and
    depths far outside int32, against the oracle, for every register-tile tuning point, and the
    same grid through the general kernels (KSA_SWEEP_GENERAL=1)."""
    torch = torch_mod()
    rng = np.random.default_rng(99)
    T, N, A = 97, 77, 2
    special = np.array([2**31 - 1, 2**31, 2**31 + 1, -(2**31), -(2**31) - 1, -(2**31) + 1, 2**40, -(2**40),
                        2**62, -(2**62), 0, 1, -1, 100, 10], dtype=np.int64)
    d_host = rng.integers(-50, 400, size=(T, A, N), dtype=np.int64)
    pick = rng.integers(0, 3, size=(T, A, N)) == 0
    d_host[pick] = special[rng.integers(0, special.size, size=int(pick.sum()))]
    pols = []
    for up, down in ((100, 10), (2**31 - 1, 2**31 - 2), (2**31 - 1, -(2**31)), (-(2**31) + 1, -(2**31)),
                     (1, 0), (0, -1), (300, 299), (2**31 - 1, 0)):
        for cu, cd in ((0, 0), (10 * S, 15 * S)):
            pols.append(ksa.Policy(attr_mask=3, scale_up_messages=up, scale_down_messages=down,
                                   scale_up_cool_down=cu, scale_down_cool_down=cd, max_pods=7))
    depth = torch.from_numpy(d_host).to(engine.dev)
    r0 = oracle_lib.fresh_state(5, 0, N, 1, 7)
    want = [oracle_lib.replay(p.to_c(), d_host, T, A, N, N, 0, *r0, want_full=False, want_summary=True).summary
            for p in pols]
    r0_dev = torch.from_numpy(r0[0]).to(engine.dev)
    for general in ("0", "1"):
        monkeypatch.setenv("KSA_SWEEP_GENERAL", general)
        for sv in variants_present(engine, 4, (0, 1, 2, 5, 6, 7, 9, 16, 17, 18, 19, 20, 21, 22, 23, 26, 33, 37, 40)):
            monkeypatch.setenv("KSA_SWEEP_VARIANT", str(sv))
            got = engine.replay_sweep(pols, depth, T, A, N, N, r0_dev).cpu().numpy().view(ksa.SUMMARY_DTYPE)
            got = got.reshape(len(pols), N)
            for i in range(len(pols)):
                assert got[i].tobytes() == want[i].tobytes(), f"general={general} variant={sv} policy {i}: {pols[i]}"
    # one unordered or out-of-range policy sends the whole grid through the general kernels
    monkeypatch.delenv("KSA_SWEEP_GENERAL")
    monkeypatch.delenv("KSA_SWEEP_VARIANT")
    extra = [ksa.Policy(attr_mask=3, scale_up_messages=2**31, scale_down_messages=5, max_pods=7),
             ksa.Policy(attr_mask=3, scale_up_messages=50, scale_down_messages=50, max_pods=7),
             ksa.Policy(attr_mask=3, scale_up_messages=50, scale_down_messages=-(2**31) - 1, max_pods=7)]
    for e in extra:
        grid = pols[:5] + [e]
        got = engine.replay_sweep(grid, depth, T, A, N, N, r0_dev).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(6, N)
        ref = oracle_lib.replay(e.to_c(), d_host, T, A, N, N, 0, *r0, want_full=False, want_summary=True).summary
        assert got[5].tobytes() == ref.tobytes() and got[0].tobytes() == want[0].tobytes(), str(e)


# ---- host-buffer pipeline (the e2e path) ----------------------------------------------------------

@pytest.mark.parametrize("A,N,T,chunk", [(1, 1000, 130, 32), (3, 517, 77, 77), (1, 4096, 64, 7)])
def test_host_pipeline_matches_oracle(engine, A, N, T, chunk):
    rng = np.random.default_rng(A * 100 + N)
    pol = ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=250, scale_down_messages=25,
                     scale_up_cool_down=12 * S, scale_down_cool_down=45 * S, max_pods=9)
    depth = ksa.pinned_empty((T, A, N), np.int64)
    depth[...] = rng.integers(0, 200, size=(T, A, N), dtype=np.int64)
    st = oracle_lib.fresh_state(8, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, N, 0, *st, want_summary=True)
    pipe = ksa.HostPipeline(0, N + 5, A, chunk)
    try:
        r, ju, jd = (np.array(x, copy=True) for x in st)
        replicas = ksa.pinned_empty((T, N), np.int32)
        events = ksa.pinned_empty((T, N), np.uint8)
        summary = np.empty(N, dtype=ksa.SUMMARY_DTYPE)
        pipe.replay(pol, depth, 0, r, ju, jd, replicas, events, summary)
        assert np.array_equal(replicas, ref.replicas) and np.array_equal(events, ref.events)
        assert summary.tobytes() == ref.summary.tobytes()
        assert np.array_equal(r, ref.state_replicas) and np.array_equal(ju, ref.state_last_up)
        assert np.array_equal(jd, ref.state_last_down)
        # summary-only call (no per-tick outputs) also advances the state
        r2, ju2, jd2 = (np.array(x, copy=True) for x in st)
        pipe.replay(pol, depth, 0, r2, ju2, jd2, None, None, summary)
        assert summary.tobytes() == ref.summary.tobytes() and np.array_equal(r2, ref.state_replicas)
    finally:
        pipe.close()


# ---- error behaviour: status codes, never an abort ------------------------------------------------

def test_errors_are_reported_not_raised_as_crashes(engine):
    torch = torch_mod()
    depth = torch.zeros((4, 1, 16), dtype=torch.int64, device=engine.dev)
    state = engine.new_state(16)
    with pytest.raises(ksa.KsaError, match="poll_period_ns"):
        engine.replay_full(ksa.Policy(attr_mask=1, poll_period=0), depth, 4, 1, 16, 16, 0, state)
    with pytest.raises(ksa.KsaError, match="attr_mask"):
        engine.replay_full(ksa.Policy(attr_mask=7), depth, 4, 1, 16, 16, 0, state)
    with pytest.raises(ksa.KsaError, match="ld="):
        engine.replay_full(ksa.Policy(attr_mask=1), depth, 4, 1, 16, 8, 0, state)
    with pytest.raises(ksa.KsaError, match="2\\^30"):
        engine.replay_full(ksa.Policy(attr_mask=1), depth, 2**30 + 1, 1, 16, 16, 0, state)
    with pytest.raises(ksa.KsaError, match="variant"):
        engine.set_variant(999, 0)
    # and the engine still works afterwards
    engine.replay_full(ksa.Policy(attr_mask=1), depth, 4, 1, 16, 16, 0, state)
    torch.cuda.synchronize()


# ---- full BASELINE size, checked through size-independent properties ------------------------------

def test_config2_full_size_properties(engine):
    """1M traces x 10k ticks, default policy, single attribute (BASELINE.json configs[1]); 136 GB.
    Checks: (a) a 2048-trace column block equals the oracle bit-for-bit over all 10k ticks,
    (b) the fingerprint of the full-mode trajectory equals the summary-mode fingerprint for every
    one of the 2^20 traces, (c) replaying again from the same state reproduces the same bytes
    (idempotence), (d) chunked == one-shot on the carry state."""
    torch = torch_mod()
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    N, T = 1 << 20, 10_000
    need = N * T * 13 + (2 << 30)
    if free < need:
        pytest.skip(f"needs {need >> 30} GiB of free device memory, have {free >> 30}")
    pol = ksa.Policy(attr_mask=1)
    gen = ksa.GenParams(seed=0x5EED0001)
    depth = engine.gen_traces(gen, 0, T, 1, 0, N, N)
    state = engine.new_state(N, seed=gen.seed, lo=1, hi=5)
    s0 = state.clone()
    replicas, events = engine.replay_full(pol, depth, T, 1, N, N, 0, state)
    # (b)
    st_s = s0.clone()
    summ = engine.replay_summary(pol, depth, T, 1, N, N, 0, st_s)
    h_full = engine.traj_hash(replicas, events, T, N, N)
    sv = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
    assert np.array_equal(h_full.cpu().numpy().view(np.uint64), sv["hash"])
    assert np.array_equal(sv["final_replicas"], state.replicas.cpu().numpy())
    assert (sv["poll_failed"] == 0).all() and (sv["up_failed"] == 0).all() and (sv["down_failed"] == 0).all()
    assert torch.equal(st_s.last_up, state.last_up) and torch.equal(st_s.last_down, state.last_down)
    # (a)
    c0, c1 = 517 * 1024, 517 * 1024 + 2048
    d_blk = depth[:, :, c0:c1].contiguous().cpu().numpy()
    ref = oracle_lib.replay(pol.to_c(), d_blk, T, 1, c1 - c0, c1 - c0, 0,
                            *oracle_lib.fresh_state(gen.seed, c0, c1 - c0, 1, 5), want_summary=True)
    assert np.array_equal(replicas[:, c0:c1].cpu().numpy(), ref.replicas)
    assert np.array_equal(events[:, c0:c1].cpu().numpy(), ref.events)
    assert sv[c0:c1].tobytes() == ref.summary.tobytes()
    # (c) idempotence: same inputs, same state -> same fingerprint
    h_first = h_full.clone()
    state2 = s0.clone()
    engine.replay_full(pol, depth, T, 1, N, N, 0, state2, replicas=replicas, events=events)
    assert torch.equal(engine.traj_hash(replicas, events, T, N, N), h_first)
    # (d) chunked replay of the same block reaches the same carry state
    state3 = s0.clone()
    for t0 in range(0, T, 2500):
        engine.replay_full(pol, depth[t0:t0 + 2500], 2500, 1, N, N, t0, state3,
                           replicas=replicas[t0:t0 + 2500], events=events[t0:t0 + 2500])
    assert torch.equal(state3.replicas, state.replicas) and torch.equal(state3.last_up, state.last_up)
    assert torch.equal(engine.traj_hash(replicas, events, T, N, N), h_first)
    del depth, replicas, events
    torch.cuda.empty_cache()


# ---- the C++ host CLI ------------------------------------------------------------------------------

@pytest.mark.parametrize("mode", ["summary", "full"])
def test_cli_totals_match_the_python_host(engine, mode):
    """kube_sqs_autoscaler_b200/ksa-replay with reference flags: its event totals equal the totals of
    the same replay driven from Python (same generator, seed, policy), which equals the oracle."""
    import os
    import re
    import subprocess

    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    N, T = 5000, 300
    r = subprocess.run([cli, "--poll-period=5s", "--scale-up-cool-down=12s", "--scale-down-cool-down=45s",
                        "--scale-up-messages=250", "--scale-down-messages=25", "--scale-up-pods=3",
                        "--scale-down-pods=2", "--max-pods=50", "--min-pods=2", "--aws-region=us-west-2",
                        "--sqs-queue-url=https://example", "--kubernetes-deployment=deploy",
                        "--traces", str(N), "--ticks", str(T), "--mode", mode, "--seed", "77"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    m = re.search(r"scale up successful (\d+) \| scale down successful (\d+) \| at max (\d+) \| at min (\d+) \| "
                  r"up cool-down skips (\d+) \| down cool-down skips (\d+)", r.stdout)
    assert m, r.stdout
    got = [int(x) for x in m.groups()]
    digest = int(re.search(r"fingerprints ([0-9a-f]{16})", r.stdout).group(1), 16)
    pol = ksa.Policy(scale_up_messages=250, scale_down_messages=25, scale_up_pods=3, scale_down_pods=2,
                     max_pods=50, min_pods=2, scale_up_cool_down=12 * S, scale_down_cool_down=45 * S)
    gen = ksa.GenParams(seed=77, low_max=25, high_min=250, high_max=1000)
    ld = (N + 15) // 16 * 16
    d_host = oracle_lib.gen_traces(gen.to_c(), 0, T, 3, 0, N, ld)
    ref = oracle_lib.replay(pol.to_c(), d_host, T, 3, N, ld, 0, *oracle_lib.fresh_state(77, 0, N, 2, 50),
                            want_full=False, want_summary=True)
    want = [int(ref.summary[k].sum()) for k in ("up_applied", "down_applied", "up_noop", "down_noop", "up_cool", "down_cool")]
    assert got == want
    assert digest == int(np.bitwise_xor.reduce(ref.summary["hash"]))
    assert "on 1 GPU(s)" in r.stdout


def test_cli_shards_over_gpus_without_changing_results(engine):
    """--gpus G: one host thread per GPU, contiguous trace shards, host gather. The totals and the
    fingerprint digest do not depend on G (global trace ids in the generator)."""
    import os
    import re
    import subprocess

    import torch

    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    args = ["--traces", "70001", "--ticks", "200", "--seed", "9", "--attribute-names=ApproximateNumberOfMessages"]
    outs = []
    for g in sorted({1, min(2, torch.cuda.device_count()), torch.cuda.device_count()}):
        r = subprocess.run([cli, *args, "--gpus", str(g)], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        outs.append((re.search(r"scale up successful.*", r.stdout).group(0),
                     re.search(r"fingerprints ([0-9a-f]{16})", r.stdout).group(1)))
    assert all(o == outs[0] for o in outs), outs
    r = subprocess.run([cli, *args, "--gpus", str(torch.cuda.device_count() + 1)], capture_output=True, text=True)
    assert r.returncode == 2 and "CUDA devices" in r.stderr


@pytest.mark.parametrize("P,N", [(1, 1), (7, 37), (9, 515), (33, 4099), (300, 700), (129, 9500), (70, 9472)])
def test_sweep_reduce_matches_restatement(engine, P, N):
    """ksa_sweep_reduce (extension): totals per policy, argmin per trace and its cost, bit-exact
    against oracle/policy_py.sweep_reduce on real sweep summaries and on wrapping random records;
    the policy totals accumulate over calls (shards add up)."""
    import torch
    from oracle import policy_py

    rng = np.random.default_rng(P * 1000 + N)
    T, S = 48, 10**9
    pols = [ksa.Policy(attr_mask=1, scale_up_messages=40 + 9 * i, scale_down_messages=3 + 2 * (i % 7),
                       scale_up_cool_down=(i % 3) * 10 * S, scale_down_cool_down=(i % 4) * 10 * S,
                       max_pods=2 + i % 9) for i in range(P)]
    depth = torch.from_numpy(rng.integers(0, 200, size=(T, 1, N), dtype=np.int64)).to(engine.dev)
    r0 = torch.ones(N, dtype=torch.int32, device=engine.dev)
    summ = engine.replay_sweep(pols, depth, T, 1, N, N, r0)
    host = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
    w = dict(w_pod_ticks=3, w_saturated=40, w_churn=11, w_blocked=-1)
    pc, bp, bc = engine.sweep_reduce(summ, P, N, **w)
    rpc, rbp, rbc = policy_py.sweep_reduce(host, **w)
    assert np.array_equal(pc.cpu().numpy(), rpc)
    assert np.array_equal(bp.cpu().numpy(), rbp) and np.array_equal(bc.cpu().numpy(), rbc)
    pc2, _, _ = engine.sweep_reduce(summ, P, N, policy_cost=pc, **w)  # accumulates
    with np.errstate(over="ignore"):
        assert np.array_equal(pc2.cpu().numpy(), rpc + rpc)
    # wrapping arithmetic and ties on synthetic records
    fake = np.zeros((P, N), dtype=ksa.SUMMARY_DTYPE)
    for f in ("up_applied", "down_applied", "up_noop", "up_cool", "down_cool"):
        fake[f] = rng.integers(0, 2**32, size=(P, N), dtype=np.uint64).astype(np.uint32)
    fake["sum_replicas"] = rng.integers(-2**62, 2**62, size=(P, N), dtype=np.int64)
    if P > 2:
        fake[P - 1] = fake[1]
    w = dict(w_pod_ticks=2**33 + 1, w_saturated=-5, w_churn=2**40, w_blocked=3)
    dev = torch.from_numpy(fake.view(np.uint8).reshape(P, N, 64)).to(engine.dev)
    pc, bp, bc = engine.sweep_reduce(dev, P, N, **w)
    rpc, rbp, rbc = policy_py.sweep_reduce(fake, **w)
    assert np.array_equal(pc.cpu().numpy(), rpc)
    assert np.array_equal(bp.cpu().numpy(), rbp) and np.array_equal(bc.cpu().numpy(), rbc)


def test_cli_sweep_mode_matches_engine_and_restatement(engine):
    """ksa-replay --mode sweep: comma lists on the reference's flags -> policy grid -> ksa_replay_sweep
    + ksa_sweep_reduce. The printed ranking (policy index, total cost, traces won) equals the one
    computed from the oracle's per-policy replays and oracle/policy_py.sweep_reduce on the same
    generated traces; chunked over traces and sharded over GPUs it stays the same."""
    import os
    import re
    import subprocess
    from oracle import policy_py

    torch = torch_mod()
    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    N, T = 700, 120
    args = [cli, "--mode", "sweep", "--poll-period=5s", "--scale-up-messages=60,100,200", "--scale-down-messages=5,20",
            "--scale-up-cool-down=0s,20s", "--scale-down-cool-down=30s", "--max-pods=4,9", "--traces", str(N),
            "--ticks", str(T), "--seed", "31", "--cost-pod-ticks", "2", "--cost-saturated", "30", "--cost-churn", "3",
            "--cost-blocked", "1", "--top", "24"]
    r = subprocess.run(args, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    rows = re.findall(r"#(\d+) policy (\d+) cost (-?\d+) best-for (\d+) traces: (.*)", r.stdout)
    assert len(rows) == 24, r.stdout
    # the same grid in the CLI's order (last tunable fastest), the same traces
    pols = [ksa.Policy(attr_mask=1, poll_period=5 * S, scale_up_messages=um, scale_down_messages=dm,
                       scale_up_cool_down=uc * S, scale_down_cool_down=30 * S, max_pods=mx)
            for um in (60, 100, 200) for dm in (5, 20) for uc in (0, 20) for mx in (4, 9)]
    gen = ksa.GenParams(seed=31, family=ksa.GEN_WALK, low_max=0, high_min=1, high_max=4 * 60,
                        walk_step=max(1, 4 * 60 // 32))
    ld = (N + 15) // 16 * 16
    walk = oracle_lib.gen_walk_init(gen.to_c(), 0, N)
    d_host = oracle_lib.gen_traces(gen.to_c(), 0, T, 1, 0, N, ld, walk=walk)
    st = oracle_lib.fresh_state(31, 0, N, 1, 4)
    summ = np.stack([oracle_lib.replay(p.to_c(), d_host, T, 1, N, ld, 0, *st, want_full=False, want_summary=True).summary
                     for p in pols])
    cost, best, _ = policy_py.sweep_reduce(summ, w_pod_ticks=2, w_saturated=30, w_churn=3, w_blocked=1)
    order = sorted(range(len(pols)), key=lambda p: (int(cost[p]), p))
    wins = np.bincount(best, minlength=len(pols))
    for k, (rank, pid, c, w, flags) in enumerate(rows):
        assert int(rank) == k + 1 and int(pid) == order[k] and int(c) == int(cost[order[k]]) and int(w) == wins[order[k]], (k, rows[k])
    p0 = pols[order[0]]
    assert f"--scale-up-messages={p0.scale_up_messages} " in rows[0][4] and f"--max-pods={p0.max_pods} " in rows[0][4]
    if torch.cuda.device_count() >= 2:
        r2 = subprocess.run(args + ["--gpus", "2"], capture_output=True, text=True, timeout=120)
        assert r2.returncode == 0, r2.stderr
        assert re.findall(r"#(\d+) policy (\d+) cost (-?\d+) best-for (\d+)", r2.stdout) == [x[:4] for x in rows]


def test_summary_counters_survive_long_saturated_runs(engine, monkeypatch):
    """The summary kernels count events in packed 5-bit fields, folded every <= 31 ticks into 10-bit
    fields, flushed every <= 30 folds. Traces on which the same event fires on EVERY tick (the worst
    case for each field) over T = 2100 ticks (more than two flushes), for every summary variant,
    both sweep kernels and the failure-mask kernel, against the oracle."""
    torch = torch_mod()
    T, N, A = 2100, 96, 1
    d_host = np.zeros((T, A, N), dtype=np.int64)
    d_host[:, 0, 0:16] = 500          # above the up threshold on every tick
    d_host[:, 0, 16:32] = 0           # below the down threshold on every tick
    d_host[:, 0, 32:48] = 50          # both thresholds at once for the overlapping policy
    rng = np.random.default_rng(3)
    d_host[:, 0, 48:] = rng.integers(0, 600, size=(T, N - 48))
    pols = [ksa.Policy(attr_mask=1, scale_up_messages=100, scale_down_messages=10, scale_up_cool_down=0,
                       scale_down_cool_down=0, max_pods=3),                       # noop on every tick
            ksa.Policy(attr_mask=1, scale_up_messages=100, scale_down_messages=10, scale_up_cool_down=10**6 * S,
                       scale_down_cool_down=10**6 * S, max_pods=3),               # cool-down skip on every tick
            ksa.Policy(attr_mask=1, scale_up_messages=40, scale_down_messages=60, scale_up_cool_down=0,
                       scale_down_cool_down=0, max_pods=2**31 - 1, min_pods=-(2**31)),  # up and down applied on every tick
            ksa.Policy(attr_mask=1, scale_up_messages=100, scale_down_messages=10, scale_up_cool_down=7 * S,
                       scale_down_cool_down=11 * S, max_pods=9)]
    depth = torch.from_numpy(d_host).to(engine.dev)
    st = oracle_lib.fresh_state(9, 0, N, 1, 3)
    want = [oracle_lib.replay(p.to_c(), d_host, T, A, N, N, 0, *st, want_full=False, want_summary=True).summary
            for p in pols]
    assert int(want[0]["up_noop"].max()) >= T - 3 and int(want[1]["down_cool"].max()) == T
    assert int(want[2]["up_applied"].max()) == T and int(want[2]["down_applied"].max()) == T
    try:
        for sv in summary_variants(engine):
            engine.set_variant(0, sv)
            for i, pol in enumerate(pols):
                got = engine.replay_summary(pol, depth, T, A, N, N, 0, engine.state_from_host(*st))
                got = got.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
                assert got.tobytes() == want[i].tobytes(), f"summary variant {sv} policy {i}"
    finally:
        engine.set_variant(0, 0)
    r0 = torch.from_numpy(st[0]).to(engine.dev)
    for general in ("0", "1"):
        monkeypatch.setenv("KSA_SWEEP_GENERAL", general)
        for sv in variants_present(engine, 4, (0, 2, 5, 9, 16, 22, 26, 27, 33, 37)):
            monkeypatch.setenv("KSA_SWEEP_VARIANT", str(sv))
            got = engine.replay_sweep(pols, depth, T, A, N, N, r0).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(len(pols), N)
            for i in range(len(pols)):
                assert got[i].tobytes() == want[i].tobytes(), f"sweep general={general} variant={sv} policy {i}"
    monkeypatch.delenv("KSA_SWEEP_GENERAL")
    monkeypatch.delenv("KSA_SWEEP_VARIANT")
    # failure mask: a failed poll / failed calls on every tick for some traces
    fail = np.zeros((T, N), dtype=np.uint8)
    fail[:, 0:8] = 1
    fail[:, 8:16] = 2
    fail[:, 16:24] = 4
    fail[:, 48:] = rng.choice(np.array([0, 0, 1, 2, 4, 7], dtype=np.uint8), size=(T, N - 48))
    f_dev = torch.from_numpy(fail).to(engine.dev)
    for fv in ["0"] + [str(i) for i in variants_present(engine, 3)]:
        monkeypatch.setenv("KSA_FAULTS_VARIANT", fv)
        for i, pol in enumerate(pols):
            ref = oracle_lib.replay(pol.to_c(), d_host, T, A, N, N, 0, *st, want_full=False, want_summary=True, fail=fail).summary
            got = engine.replay_summary(pol, depth, T, A, N, N, 0, engine.state_from_host(*st), fail=f_dev)
            assert got.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.tobytes(), f"faults variant {fv} policy {i}"
    assert int(ref["poll_failed"].max()) == T


def test_rendered_log_of_a_gpu_trajectory_matches_the_restated_reference(engine):
    """a7 end to end: the event bytes and replica counts written by ksa_replay_full_faults, turned
    back into log messages (render_log), are the lines the restated reference objects print."""
    from oracle import policy_py

    torch = torch_mod()
    rng = np.random.default_rng(21)
    T, N = 60, 48
    pol = ksa.Policy(attr_mask=3, scale_up_messages=90, scale_down_messages=95, scale_up_cool_down=10 * S,
                     scale_down_cool_down=15 * S, scale_up_pods=2, max_pods=6)  # overlapping thresholds
    d_host = rng.integers(0, 100, size=(T, 2, N), dtype=np.int64)
    fail = rng.choice(np.array([0, 0, 0, 0, 1, 2, 4], dtype=np.uint8), size=(T, N))
    st = oracle_lib.fresh_state(4, 0, N, 1, 6)
    replicas, events = engine.replay_full(pol, torch.from_numpy(d_host).to(engine.dev), T, 2, N, N, 0,
                                          engine.state_from_host(*st), fail=torch.from_numpy(fail).to(engine.dev))
    replicas, events = replicas.cpu().numpy(), events.cpu().numpy()
    for n in range(N):
        text = []
        rows = [tuple(int(x) for x in d_host[t, :, n]) for t in range(T)]
        policy_py.replay_trace(pol.as_abi_dict(), rows, int(st[0][n]), fail_rows=[int(f) for f in fail[:, n]],
                               text_log=text)
        got = ksa.render_log(pol, int(st[0][n]), d_host[:, :, n].sum(axis=1), replicas[:, n], events[:, n])
        assert got == text, f"trace {n}"


@pytest.mark.parametrize("A", [1, 3])
def test_sweep_random_grids_mixed_tiles(engine, A):
    """Random policy grids (ordered, overlapping and out-of-int32 thresholds mixed, P not a multiple
    of the tile) on adversarial depths: every (policy, trace) summary equals the oracle's. Exercises
    the split launch with arbitrary tile patterns."""
    torch = torch_mod()
    rng = np.random.default_rng(4242 + A)
    for case in range(6):
        P = int(rng.integers(1, 24))
        T = int(rng.integers(1, 80))
        N = int(rng.choice([1, 33, 130, 257]))
        mask = (1 << A) - 1
        pols = []
        for _ in range(P):
            p = random_policy(rng, A)
            kind = rng.integers(0, 3)
            if kind == 0:      # an everyday ordered policy
                p = ksa.Policy(attr_mask=mask, scale_up_messages=int(rng.integers(50, 400)),
                               scale_down_messages=int(rng.integers(0, 49)), poll_period=p.poll_period,
                               scale_up_cool_down=p.scale_up_cool_down, scale_down_cool_down=p.scale_down_cool_down,
                               max_pods=int(rng.integers(2, 12)))
            else:              # anything, including thresholds beyond int32 and overlaps
                p = ksa.Policy(attr_mask=mask, poll_period=p.poll_period, scale_up_cool_down=p.scale_up_cool_down,
                               scale_down_cool_down=p.scale_down_cool_down, scale_up_messages=p.scale_up_messages,
                               scale_down_messages=p.scale_down_messages, scale_up_pods=p.scale_up_pods,
                               scale_down_pods=p.scale_down_pods, max_pods=p.max_pods, min_pods=p.min_pods)
            pols.append(p)
        depth_np = adversarial_depth(rng, T, A, N, pols[0])
        r0 = rng.integers(-2, 9, size=N).astype(np.int32)
        st = (r0, np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
        got = engine.replay_sweep(pols, torch.from_numpy(depth_np).to(engine.dev), T, A, N, N,
                                  torch.from_numpy(r0).to(engine.dev))
        got = got.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
        for i, pol in enumerate(pols):
            ref = oracle_lib.replay(pol.to_c(), depth_np, T, A, N, N, 0, *st, want_full=False, want_summary=True)
            assert got[i].tobytes() == ref.summary.tobytes(), f"A={A} case {case} policy {i}/{P}: {pol}"


@pytest.mark.parametrize("A", [1, 3])
def test_sweep_shared_threshold_tiles(engine, A, monkeypatch):
    """Grids laid out as real parameter grids are (thresholds vary slowest): runs of policies with the
    same two thresholds, so that whole register tiles share them and take the shared-threshold kernel
    (one pair of threshold tests per (trace, tick) for the tile). Mixed with tiles that straddle a
    threshold change (narrow kernel), tiles with overlapping thresholds and thresholds beyond int32
    (general kernel; a shared tile need not fit int32), depths at the thresholds and far outside
    int32, run lengths that are not multiples of the tile, and replica counts large enough to leave
    the 32-bit window sum of the summary. Every (policy, trace) summary equals the oracle's, with the
    three-way split, with shared tiles not recognised (KSA_SWEEP_NOSHARED) and through the general
    kernels alone."""
    torch = torch_mod()
    rng = np.random.default_rng(777 + A)
    mask = (1 << A) - 1
    for case in range(4):
        thresholds = [(100, 10), (250, 25), (2**31 + 5, -(2**31) - 5), (7, 6), (2**31 - 1, -(2**31)), (50, 60),
                      (2**40, 2**40 - 1)]
        pols = []
        for up, down in (thresholds[i] for i in rng.permutation(len(thresholds))[:5]):
            for _ in range(int(rng.choice([3, 4, 5, 8, 9]))):  # run of policies sharing the thresholds
                big = rng.integers(0, 4) == 0
                pols.append(ksa.Policy(attr_mask=mask, scale_up_messages=up, scale_down_messages=down,
                                       poll_period=5 * S, scale_up_cool_down=int(rng.integers(0, 7)) * 5 * S,
                                       scale_down_cool_down=int(rng.integers(0, 9)) * 3 * S,
                                       max_pods=int(rng.integers(2, 12)) if not big else 2**31 - 1,
                                       min_pods=int(rng.integers(0, 3)),
                                       scale_up_pods=int(rng.integers(1, 4)) if not big else 2**27 + 11,
                                       scale_down_pods=int(rng.integers(1, 4)) if not big else 2**26 + 3))
        P = len(pols)
        T = int(rng.choice([1, 2, 3, 40, 75, 131]))
        N = int(rng.choice([1, 33, 130]))
        depth_np = adversarial_depth(rng, T, A, N, pols[0])
        edge = np.array([100, 99, 10, 11, 250, 25, 7, 6, 2**31 + 5, 2**31 + 4, -(2**31) - 5, -(2**31) - 4, 2**31 - 1,
                         -(2**31), 2**40, 2**40 - 1, 50, 60], dtype=np.int64)
        pick = rng.integers(0, 2, size=depth_np.shape) == 0
        depth_np[pick] = edge[rng.integers(0, edge.size, size=int(pick.sum()))] // (A if A > 1 else 1)
        r0 = rng.integers(-2, 9, size=N).astype(np.int32)
        st = (r0, np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
        want = [oracle_lib.replay(p.to_c(), depth_np, T, A, N, N, 0, *st, want_full=False, want_summary=True).summary
                for p in pols]
        d = torch.from_numpy(depth_np).to(engine.dev)
        r0d = torch.from_numpy(r0).to(engine.dev)
        for env in ({}, {"KSA_SWEEP_NOSHARED": "1"}, {"KSA_SWEEP_GENERAL": "1"}):
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            got = engine.replay_sweep(pols, d, T, A, N, N, r0d).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
            for k in env:
                monkeypatch.delenv(k)
            for i in range(P):
                assert got[i].tobytes() == want[i].tobytes(), f"A={A} case {case} env {env} policy {i}/{P}: {pols[i]}"


@pytest.mark.parametrize("A", [1, 2])
def test_summary_sum_is_exact_beyond_the_32_bit_window(engine, A):
    """The summary kernels add the replica counts of up to 31 ticks in 32 bits and fold; a trace whose
    replica count leaves +-2^26 gets its sum recomputed exactly. Traces far beyond that (steps of
    2^27, limits at the int32 extremes, negative and wrapping counts, Q6) next to ordinary ones, in
    chunks (the running minimum / maximum come from the record) and with a failure mask: the
    64-bit sum_replicas and everything else equal the oracle's, for every summary variant."""
    torch = torch_mod()
    rng = np.random.default_rng(31337 + A)
    mask = (1 << A) - 1
    N, T = 97, 300
    pols = [ksa.Policy(attr_mask=mask, scale_up_messages=100, scale_down_messages=10, scale_up_cool_down=0,
                       scale_down_cool_down=5 * S, max_pods=2**31 - 1, min_pods=-(2**31), scale_up_pods=2**27 + 1,
                       scale_down_pods=2**25 + 7),
            ksa.Policy(attr_mask=mask, scale_up_messages=10, scale_down_messages=100, scale_up_cool_down=10 * S,
                       scale_down_cool_down=0, max_pods=2**31 - 1, min_pods=-(2**31), scale_up_pods=2**30,
                       scale_down_pods=2**29 + 5),
            ksa.Policy(attr_mask=mask, scale_up_messages=100, scale_down_messages=10, max_pods=2**26 + 40,
                       min_pods=2**26 - 40, scale_up_pods=7, scale_down_pods=9, scale_up_cool_down=0,
                       scale_down_cool_down=0)]
    depth_np = rng.integers(0, 140 // A, size=(T, A, N), dtype=np.int64)
    fail_np = rng.choice(np.array([0, 0, 0, 0, 1, 2, 4], dtype=np.uint8), size=(T, N))
    r0 = rng.integers(-5, 9, size=N).astype(np.int32)
    r0[::7] = 2**26 - 3          # starts next to the limit of the fast path
    r0[3::11] = -(2**31) + 5
    r0[5::13] = 2**31 - 9
    st = (r0, np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
    d = torch.from_numpy(depth_np).to(engine.dev)
    f = torch.from_numpy(fail_np).to(engine.dev)
    try:
        for pol in pols:
            for fail_arr, fail_dev in ((None, None), (fail_np, f)):
                ref = oracle_lib.replay(pol.to_c(), depth_np, T, A, N, N, 0, *st, want_full=False, want_summary=True,
                                        fail=fail_arr)
                ids = summary_variants(engine) if fail_dev is None else [0] + variants_present(engine, 3)
                for sv in ids:
                    if fail_dev is None:
                        engine.set_variant(0, sv)
                    else:
                        os.environ["KSA_FAULTS_VARIANT"] = str(sv)
                        engine.set_variant(0, -2 if sv == 0 else 0)
                    for chunks in ((T,), (64, 31, 1, T - 96)):
                        state = engine.state_from_host(*st)
                        summ = engine.new_summary(N)
                        t0 = 0
                        for tc in chunks:
                            engine.replay_summary(pol, d[t0:t0 + tc].contiguous(), tc, A, N, N, t0, state, summary=summ,
                                                  fail=None if fail_dev is None else fail_dev[t0:t0 + tc].contiguous())
                            t0 += tc
                        got = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
                        assert got.tobytes() == ref.summary.tobytes(), f"A={A} variant {sv} chunks {chunks} faults {fail_dev is not None}: {pol}"
        # the sweep keeps the same accumulators
        r0d = torch.from_numpy(r0).to(engine.dev)
        sweep_pols = [p for p in pols for _ in range(3)]
        got = engine.replay_sweep(sweep_pols, d, T, A, N, N, r0d).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(len(sweep_pols), N)
        for i, pol in enumerate(sweep_pols):
            ref = oracle_lib.replay(pol.to_c(), depth_np, T, A, N, N, 0, *st, want_full=False, want_summary=True)
            assert got[i].tobytes() == ref.summary.tobytes(), f"sweep policy {i}: {pol}"
    finally:
        os.environ.pop("KSA_FAULTS_VARIANT", None)
        engine.set_variant(0, 0)


def test_registered_host_memory_takes_the_summaries(engine):
    """ksa_host_register page-locks memory the caller already has (a slice of a shared mapping in the
    multi-GPU gather of bench.py): summaries copied into a registered slice of an anonymous mapping
    must arrive exactly, neighbours untouched; registering nothing is an error."""
    import ctypes as C
    import mmap

    import torch

    from kube_sqs_autoscaler_b200 import _abi

    N, T, page = 512, 64, mmap.PAGESIZE
    rng = np.random.default_rng(21)
    pol = ksa.Policy(attr_mask=1)
    depth = rng.integers(0, 300, size=(T, 1, N), dtype=np.int64)
    st = oracle_lib.fresh_state(3, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth, T, 1, N, N, 0, *st, want_summary=True)
    summ = engine.replay_summary(pol, torch.from_numpy(depth).to(engine.dev), T, 1, N, N, 0, engine.state_from_host(*st))
    m = mmap.mmap(-1, 3 * N * 64)  # three slices; the middle one (page-aligned: N * 64 = 8 pages) is ours
    assert (N * 64) % page == 0
    buf = np.frombuffer(m, dtype=np.uint8)
    buf[:] = 0xEE
    mid = buf.ctypes.data + N * 64
    _abi.check(engine.lib.ksa_host_register(mid, N * 64))
    try:
        _abi.check(engine.lib.ksa_memcpy_d2h(mid, summ.data_ptr(), N * 64, torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
    finally:
        _abi.check(engine.lib.ksa_host_unregister(mid))
    assert buf[N * 64:2 * N * 64].tobytes() == ref.summary.tobytes()
    assert (buf[:N * 64] == 0xEE).all() and (buf[2 * N * 64:] == 0xEE).all()
    assert engine.lib.ksa_host_register(None, 0) == _abi.KSA_EINVAL
    del buf
    m.close()


def test_random_cases_over_every_other_path(engine):
    """A short run of tools/fuzz_paths.py: the plain kernels in every compiled variant (masks, T-chunks
    handing the carry state on), policy grids through both sweep entry points in every tile mode and on
    into the reduction, and the host pipeline with random stage lengths; start counts anywhere in int32."""
    from tools import fuzz_paths

    assert fuzz_paths.run(200, 21, engine) == 0
