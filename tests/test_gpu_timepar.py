"""Full-trajectory replay with the time axis cut into chunks (csrc/ksa_timepar.cu): the path
ksa_replay_full takes by itself when the traces are too few to fill the GPU. It must give exactly
what the sequential replay gives whatever the number of chunks and whether or not the speculative
chunks ever converge to the true trajectory, so the chunk count is forced here
(KSA_TIMEPAR_CHUNKS) on small cases: random adversarial policies, and traces built never to
converge. Everything against the CPU oracle, through the C ABI, bit for bit."""
from __future__ import annotations

import os

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from tests import oracle_lib
from tests.helpers import S
from tests.test_gpu_parity import adversarial_depth, random_policy, random_state

pytestmark = pytest.mark.gpu


TP_LAUNCHES = 4  # pass 1, 2a, 2b, 3


def chunks_used(T, chunks):
    """Chunks the library makes of a forced count: lengths are rounded up to 8 ticks."""
    if chunks < 2 or (T + chunks - 1) // chunks < 16:
        return 0
    L = ((T + chunks - 1) // chunks + 7) // 8 * 8
    return (T + L - 1) // L


def run_and_check(engine, pol, depth, T, A, N, ld, tick0, st, chunks, label, warm=None, wide=None):
    """Replay with `chunks` forced (None = the library's own choice); returns the launches it took."""
    import torch

    ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, tick0, *st)
    d = torch.from_numpy(np.ascontiguousarray(depth)).to(engine.dev)
    for key, val in (("KSA_TIMEPAR_CHUNKS", chunks), ("KSA_TIMEPAR_WARM", warm), ("KSA_TIMEPAR_WIDE", wide)):
        if val is None:
            os.environ.pop(key, None)
        else:
            os.environ[key] = str(val)
    try:
        state = engine.state_from_host(*st)
        engine.launch_count()
        r, e = engine.replay_full(pol, d, T, A, N, ld, tick0, state)
        launches = engine.launch_count()
        r, e = r.cpu().numpy(), e.cpu().numpy()
    finally:
        for key in ("KSA_TIMEPAR_CHUNKS", "KSA_TIMEPAR_WARM", "KSA_TIMEPAR_WIDE"):
            os.environ.pop(key, None)
    where = np.argwhere(r[:, :N] != ref.replicas[:, :N])
    assert where.size == 0, f"{label}: replicas differ first at (tick, trace) {where[0]} of {where.shape[0]}"
    assert np.array_equal(e[:, :N], ref.events[:, :N]), f"{label}: events differ"
    assert np.array_equal(state.replicas.cpu().numpy(), ref.state_replicas), f"{label}: carry replicas differ"
    assert np.array_equal(state.last_up.cpu().numpy(), ref.state_last_up), f"{label}: last_up differs"
    assert np.array_equal(state.last_down.cpu().numpy(), ref.state_last_down), f"{label}: last_down differs"
    return launches


@pytest.mark.parametrize("A", [1, 2, 3])
def test_random_policies_any_chunk_count(engine, A):
    rng = np.random.default_rng(5150 + A)
    for case in range(10):
        pol = random_policy(rng, A)
        T = int(rng.choice([40, 97, 260, 700]))
        N = int(rng.choice([1, 33, 130, 1000, 64, 256]))  # (multiples of 8 on a padded pitch: 8 traces per thread)
        ld = (N + 15) // 16 * 16 if case % 2 else N
        tick0 = int(rng.choice([0, 0, 23, 10**6, 2**40]))
        depth = adversarial_depth(rng, T, A, ld, pol)
        st = random_state(rng, N, tick0)
        for chunks in (2, 3, 5, 16):
            for warm, wide in ((None, None), (0, 1), (8, 0), (40, 1)):
                took = run_and_check(engine, pol, depth, T, A, N, ld, tick0, st, chunks,
                                     f"A={A} case {case} T={T} N={N} ld={ld} chunks={chunks} warm={warm} wide={wide} {pol}",
                                     warm=warm, wide=wide)
                if chunks_used(T, chunks) >= 2:
                    assert took == TP_LAUNCHES, "the forced chunk count was not used"


def test_traces_that_never_converge(engine):
    """The speculative chunks start from the trace's initial replica count with both gates open.
    Cases where that never meets the true trajectory: replicas outside [min, max] with the one-sided
    clamps (Q4: an r0 above max only comes down by down-events), min > max, cool-downs longer than
    the whole call (the true gates stay shut, the guessed ones are open), a constant depth over the
    threshold (scale calls every cool-down period, the phase set by history), and wrapping steps."""
    rng = np.random.default_rng(99)
    T = 400
    for N in (77, 80):  # one trace per thread / eight
        _never_converging_cases(engine, rng, T, N)


def _never_converging_cases(engine, rng, T, N):
    cases = [
        ("r0 outside the clamp", ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=0, max_pods=5, min_pods=1),
         rng.integers(0, 400, size=(T, 1, N)), rng.integers(-40, 60, size=N)),
        ("min above max", ksa.Policy(attr_mask=1, max_pods=2, min_pods=9, scale_up_cool_down=5 * S, scale_down_cool_down=0),
         rng.integers(0, 400, size=(T, 1, N)), rng.integers(-3, 14, size=N)),
        ("cool-downs longer than the call", ksa.Policy(attr_mask=1, scale_up_cool_down=10**6 * S, scale_down_cool_down=10**5 * S),
         rng.integers(0, 400, size=(T, 1, N)), rng.integers(1, 6, size=N)),
        ("constant load, phase from history", ksa.Policy(attr_mask=1, scale_up_cool_down=35 * S, scale_down_cool_down=55 * S,
                                                          max_pods=1000, min_pods=-1000),
         np.concatenate([np.full((T // 2, 1, N), 500), np.full((T - T // 2, 1, N), 0)]), rng.integers(1, 6, size=N)),
        ("wrapping steps", ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=5 * S, max_pods=2**31 - 1,
                                      min_pods=-(2**31), scale_up_pods=2**30 + 12345, scale_down_pods=2**29 + 7),
         rng.integers(0, 400, size=(T, 1, N)), rng.integers(-5, 9, size=N)),
    ]
    for name, pol, depth, r0 in cases:
        depth = np.ascontiguousarray(depth, dtype=np.int64)
        for tick0, ju, jd in ((0, 0, 0), (1000, 998, 1000)):
            st = (r0.astype(np.int32), np.full(N, ju, dtype=np.int64), np.full(N, jd, dtype=np.int64))
            for chunks in (2, 7, 25):
                for warm in (None, 0, 16):
                    run_and_check(engine, pol, depth, T, 1, N, N, tick0, st, chunks,
                                  f"{name} tick0={tick0} chunks={chunks} warm={warm}", warm=warm)


def test_every_checkpoint_and_the_automatic_choice(engine):
    """Chunks long enough to hold every checkpoint (64, 256, 1024 ticks), with cool-downs that
    keep the speculative and the true gates apart for hundreds of ticks, so that convergence happens
    at the late checkpoints or not at all; then the library's own choice of chunks for a small block
    (N = 4096 x T = 8192 takes the time-parallel path: four launches) and a block it must leave to
    the plain kernels (too few ticks)."""
    rng = np.random.default_rng(4)
    N, T = 64, 9000
    for cu, cd in ((10, 30), (700, 90), (3000, 2500)):
        pol = ksa.Policy(attr_mask=1, scale_up_cool_down=cu * 5 * S, scale_down_cool_down=cd * 5 * S, max_pods=40, min_pods=0)
        depth = rng.integers(0, 400, size=(T, 1, N), dtype=np.int64)
        depth[rng.integers(0, 3, size=depth.shape) == 0] = 0
        st = (rng.integers(0, 41, size=N).astype(np.int32), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
        for chunks in (2, 3):
            for warm in (None, 0):
                assert run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"cool {cu}/{cd} chunks={chunks} warm={warm}",
                                     warm=warm) == TP_LAUNCHES
    N, T = 4096, 8192
    pol = ksa.Policy(attr_mask=1)
    gen = ksa.GenParams(seed=11)
    depth = oracle_lib.gen_traces(gen.to_c(), 0, T, 1, 0, N, N)
    st = oracle_lib.fresh_state(gen.seed, 0, N, 1, 5)
    assert run_and_check(engine, pol, depth, T, 1, N, N, 0, st, None, "automatic, 4096 x 8192") == TP_LAUNCHES
    assert run_and_check(engine, pol, depth[:512], 512, 1, N, N, 0, st, None, "automatic, 4096 x 512") <= 2
