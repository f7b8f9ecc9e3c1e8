"""Replay with the time axis cut into chunks (csrc/ksa_timepar.cu): the path the four replay entry
points (full / summary, with or without a failure mask) take by themselves when the traces are too
few to fill the GPU. It must give exactly what the sequential replay gives whatever the number of
chunks and whether or not the speculative chunks ever converge to the true trajectory, so the chunk
count is forced here (KSA_TIMEPAR_CHUNKS) on small cases: random adversarial policies, and traces
built never to converge. Every case goes through the full and the summary entry point. Everything
against the CPU oracle, through the C ABI, bit for bit."""
from __future__ import annotations

import os

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from tests import oracle_lib
from tests.helpers import S
from tests.test_gpu_parity import adversarial_depth, random_policy, random_state

pytestmark = pytest.mark.gpu


TP_LAUNCHES = 10  # pass 1, 2a, 2b, 3a, the second replay, the chunks written from the closed form, 3b, the shift of replica counts (two kernels), the kernel for traces given up
TP_MASK_LAUNCHES = 9  # with a failure mask no chunk has a closed form (and a probe comes first when there are traces to sample)
TP_SUMMARY_LAUNCHES = 8  # pass 0 (entry states), 1, 2a, 2b, 3a, the second replay of listed chunks, 3b, and that kernel


def chunks_used(T, chunks):
    """Chunks the library makes of a forced count: lengths are rounded up to 8 ticks."""
    if chunks < 2 or (T + chunks - 1) // chunks < 16:
        return 0
    L = ((T + chunks - 1) // chunks + 7) // 8 * 8
    return (T + L - 1) // L


def run_and_check(engine, pol, depth, T, A, N, ld, tick0, st, chunks, label, warm=None, wide=None, fail=None, probe=None):
    """Full and summary replay with `chunks` forced (None = the library's own choice), `fail` = a
    [T][ld] failure mask for the *_faults entry points; returns the launches the full replay took."""
    import torch

    ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, tick0, *st, want_summary=True, fail=fail)
    d = torch.from_numpy(np.ascontiguousarray(depth)).to(engine.dev)
    f = None if fail is None else torch.from_numpy(np.ascontiguousarray(fail, dtype=np.uint8)).to(engine.dev)
    for key, val in (("KSA_TIMEPAR_CHUNKS", chunks), ("KSA_TIMEPAR_WARM", warm), ("KSA_TIMEPAR_WIDE", wide),
                     ("KSA_TIMEPAR_PROBE", probe)):
        if val is None:
            os.environ.pop(key, None)
        else:
            os.environ[key] = str(val)
    try:
        state = engine.state_from_host(*st)
        engine.launch_count()
        r, e = engine.replay_full(pol, d, T, A, N, ld, tick0, state, fail=f)
        launches = engine.launch_count()
        r, e = r.cpu().numpy(), e.cpu().numpy()
        state2 = engine.state_from_host(*st)
        summ = engine.new_summary(N)
        engine.launch_count()
        engine.replay_summary(pol, d, T, A, N, ld, tick0, state2, summary=summ, fail=f)
        launches_summary = engine.launch_count()
        summ = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
    finally:
        for key in ("KSA_TIMEPAR_CHUNKS", "KSA_TIMEPAR_WARM", "KSA_TIMEPAR_WIDE", "KSA_TIMEPAR_PROBE"):
            os.environ.pop(key, None)
    where = np.argwhere(r[:, :N] != ref.replicas[:, :N])
    assert where.size == 0, f"{label}: replicas differ first at (tick, trace) {where[0]} of {where.shape[0]}"
    assert np.array_equal(e[:, :N], ref.events[:, :N]), f"{label}: events differ"
    for name, s in (("full", state), ("summary", state2)):
        assert np.array_equal(s.replicas.cpu().numpy(), ref.state_replicas), f"{label}: {name}: carry replicas differ"
        assert np.array_equal(s.last_up.cpu().numpy(), ref.state_last_up), f"{label}: {name}: last_up differs"
        assert np.array_equal(s.last_down.cpu().numpy(), ref.state_last_down), f"{label}: {name}: last_down differs"
    if summ.tobytes() != ref.summary.tobytes():
        for field in ksa.SUMMARY_DTYPE.names:
            bad = np.flatnonzero(summ[field] != ref.summary[field])
            assert bad.size == 0, (f"{label}: summary field {field} differs for {bad.size} traces, first {bad[0]}: "
                                   f"{summ[field][bad[0]]} != {ref.summary[field][bad[0]]}")
    if chunks is not None and chunks_used(T, chunks) >= 2:
        assert launches_summary == TP_SUMMARY_LAUNCHES, f"{label}: the summary replay did not use the forced chunk count"
    return launches


def random_mask(rng, T, ld, p=0.08):
    """KSA_FAIL_* bits, each set with probability p."""
    return ((rng.random((T, ld)) < p) * 1 + (rng.random((T, ld)) < p) * 2 + (rng.random((T, ld)) < p) * 4).astype(np.uint8)


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
        fail = random_mask(rng, T, ld) if case % 3 == 0 else None  # the *_faults entry points
        for chunks in (2, 3, 5, 16):
            for warm, wide in ((None, None), (0, 1), (8, 0), (40, 1)):
                took = run_and_check(engine, pol, depth, T, A, N, ld, tick0, st, chunks,
                                     f"A={A} case {case} T={T} N={N} ld={ld} chunks={chunks} warm={warm} wide={wide} "
                                     f"mask={fail is not None} {pol}", warm=warm, wide=wide, fail=fail)
                if chunks_used(T, chunks) >= 2:  # (with a failure mask the probe comes first, if there are 32 traces to sample)
                    assert took == (TP_LAUNCHES if fail is None else TP_MASK_LAUNCHES + (1 if N >= 32 else 0)), "the forced chunk count was not used"


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
            run_and_check(engine, pol, depth, T, 1, N, N, tick0, st, 7, f"{name} tick0={tick0} with failures",
                          warm=16, fail=random_mask(rng, T, N, 0.2))


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
    fail = oracle_lib.gen_faults(7, 5, 4, 4, 0, T, 0, N, N)
    assert run_and_check(engine, pol, depth, T, 1, N, N, 0, st, None, "automatic, 4096 x 8192, failures",
                         fail=fail) == TP_MASK_LAUNCHES + 1  # (with a failure mask the probe comes first)
    assert run_and_check(engine, pol, depth[:512], 512, 1, N, N, 0, st, None, "automatic, 4096 x 512") <= 2


def test_summary_calls_accumulate_across_chunked_calls(engine):
    """ksa_replay_summary adds to the records it is given: two calls, each cut into chunks of its
    own, must leave what one sequential replay of all ticks leaves (counters, sum, min / max, and the
    fingerprint continued from the first call's)."""
    import torch

    rng = np.random.default_rng(77)
    for A, fail_p in ((1, None), (3, 0.1)):
        pol = random_policy(rng, A)
        T1, T2, N, tick0 = 330, 270, 200, 17
        T = T1 + T2
        depth = adversarial_depth(rng, T, A, N, pol)
        fail = None if fail_p is None else random_mask(rng, T, N, fail_p)
        st = random_state(rng, N, tick0)
        ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, N, tick0, *st, want_summary=True, fail=fail)
        d = torch.from_numpy(np.ascontiguousarray(depth)).to(engine.dev)
        f = None if fail is None else torch.from_numpy(fail).to(engine.dev)
        os.environ["KSA_TIMEPAR_CHUNKS"] = "4"
        os.environ["KSA_TIMEPAR_WARM"] = "16"
        try:
            state = engine.state_from_host(*st)
            summ = engine.new_summary(N)
            engine.launch_count()
            engine.replay_summary(pol, d[:T1], T1, A, N, N, tick0, state, summary=summ, fail=None if f is None else f[:T1])
            assert engine.launch_count() == TP_SUMMARY_LAUNCHES
            engine.replay_summary(pol, d[T1:], T2, A, N, N, tick0 + T1, state, summary=summ,
                                  fail=None if f is None else f[T1:])
        finally:
            os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
            os.environ.pop("KSA_TIMEPAR_WARM", None)
        got = summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
        assert got.tobytes() == ref.summary.tobytes(), f"A={A}: accumulated summary differs"
        assert np.array_equal(state.replicas.cpu().numpy(), ref.state_replicas)
        assert np.array_equal(state.last_up.cpu().numpy(), ref.state_last_up)
        assert np.array_equal(state.last_down.cpu().numpy(), ref.state_last_down)


def test_replica_counts_beyond_the_32_bit_window_in_chunks(engine):
    """The summary kernel's 32-bit window sum falls back to an exact recomputation when replica
    counts leave +-2^26 (ksa_summary_acc); in a chunked call that recomputation starts from the
    chunk's own entry state."""
    rng = np.random.default_rng(3)
    T, N = 320, 96
    pol = ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=2 * 5 * S, max_pods=2**31 - 1,
                     min_pods=-(2**31), scale_up_pods=2**27 + 5, scale_down_pods=2**25)
    depth = np.ascontiguousarray(rng.integers(0, 400, size=(T, 1, N)), dtype=np.int64)
    st = (rng.integers(-5, 9, size=N).astype(np.int32), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
    for chunks in (2, 5):
        run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"wide replica counts, chunks={chunks}", warm=8)


def test_chunked_paths_never_write_outside_their_outputs(engine):
    """Canaries (compute-sanitizer is closed on this GPU pool): the outputs sit inside larger buffers
    filled with a marker, with padding columns in every row; after a chunked call of each of the four
    entry points the marker must be intact before, after and beside them, and the caller's state and
    summary arrays must be the only other things written."""
    import torch

    rng = np.random.default_rng(8)
    guard = 32
    for A, N, ld, T, chunks, with_mask in ((1, 64, 80, 203, 3, False), (1, 72, 96, 417, 7, True),
                                           (3, 9, 16, 160, 4, True), (2, 130, 144, 95, 2, False)):
        pol = random_policy(rng, A)
        depth_np = adversarial_depth(rng, T, A, ld, pol)
        fail_np = random_mask(rng, T, ld) if with_mask else None
        st = random_state(rng, N, 5)
        ref = oracle_lib.replay(pol.to_c(), depth_np, T, A, N, ld, 5, *st, want_summary=True, fail=fail_np)
        depth = torch.from_numpy(np.ascontiguousarray(depth_np)).to(engine.dev)
        fail = None if fail_np is None else torch.from_numpy(fail_np).to(engine.dev)
        rbuf = torch.full((T + 2 * guard, ld), -7777, dtype=torch.int32, device=engine.dev)
        ebuf = torch.full((T + 2 * guard, ld), 0xEE, dtype=torch.uint8, device=engine.dev)
        sbuf = torch.full((N + 2 * guard, 64), 0xEE, dtype=torch.uint8, device=engine.dev)
        stbuf = [torch.full((N + 2 * guard,), 0x5A5A5A5A, dtype=dt, device=engine.dev) for dt in (torch.int32, torch.int64, torch.int64)]
        os.environ["KSA_TIMEPAR_CHUNKS"] = str(chunks)
        os.environ["KSA_TIMEPAR_WARM"] = "16"
        try:
            for mode in ("full", "summary"):
                for buf, host in zip(stbuf, st):
                    buf[guard:guard + N] = torch.from_numpy(np.asarray(host)).to(engine.dev)
                state = ksa.TraceState(*(b[guard:guard + N] for b in stbuf))
                summ = sbuf[guard:guard + N]
                engine.lib.ksa_summary_clear(summ.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
                engine.launch_count()
                if mode == "full":
                    engine.replay_full(pol, depth, T, A, N, ld, 5, state, replicas=rbuf[guard:guard + T],
                                       events=ebuf[guard:guard + T], fail=fail)
                    assert engine.launch_count() == (TP_MASK_LAUNCHES + (1 if N >= 32 else 0) if with_mask else TP_LAUNCHES)  # (the probe)
                else:
                    engine.replay_summary(pol, depth, T, A, N, ld, 5, state, summary=summ, fail=fail)
                    assert engine.launch_count() == TP_SUMMARY_LAUNCHES
                label = f"{mode} A={A} N={N} ld={ld} T={T} chunks={chunks} mask={with_mask}"
                assert (rbuf[:guard] == -7777).all() and (rbuf[guard + T:] == -7777).all(), label
                assert (ebuf[:guard] == 0xEE).all() and (ebuf[guard + T:] == 0xEE).all(), label
                assert (rbuf[guard:guard + T, N:] == -7777).all() and (ebuf[guard:guard + T, N:] == 0xEE).all(), label
                assert (sbuf[:guard] == 0xEE).all() and (sbuf[guard + N:] == 0xEE).all(), label
                for buf in stbuf:
                    assert (buf[:guard] == 0x5A5A5A5A).all() and (buf[guard + N:] == 0x5A5A5A5A).all(), label
                assert np.array_equal(stbuf[0][guard:guard + N].cpu().numpy(), ref.state_replicas), label
                assert np.array_equal(stbuf[1][guard:guard + N].cpu().numpy(), ref.state_last_up), label
                assert np.array_equal(stbuf[2][guard:guard + N].cpu().numpy(), ref.state_last_down), label
                if mode == "full":
                    assert np.array_equal(rbuf[guard:guard + T, :N].cpu().numpy(), ref.replicas[:, :N]), label
                    assert np.array_equal(ebuf[guard:guard + T, :N].cpu().numpy(), ref.events[:, :N]), label
                else:
                    assert summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.summary.tobytes(), label
        finally:
            os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
            os.environ.pop("KSA_TIMEPAR_WARM", None)


def test_traces_that_go_quiet_or_stay_away_from_the_limits(engine):
    """A chunk replayed from the right gates but a replica count that is off by an offset is not
    replayed again when it never came near a limit: pass 3 shifts its record (summary mode). The two
    situations that lead there: a trace that goes quiet after some activity (no scale call for the
    rest of the call: the state never forgets, a guess never meets it), and limits so far apart that
    every call is applied in full. Also the neighbours of that rule: counts that touch a limit in one
    of the two runs only, steps that wrap, failures in between."""
    rng = np.random.default_rng(41)
    T = 640
    for N in (70, 64):
        cases = []
        quiet = rng.integers(0, 400, size=(T, 1, N))
        quiet[90:] = 50  # between the default thresholds from tick 90 on
        cases.append(("goes quiet", ksa.Policy(attr_mask=1), quiet, rng.integers(1, 6, size=N)))
        cases.append(("goes quiet outside the limits", ksa.Policy(attr_mask=1, max_pods=3, min_pods=2), quiet,
                      rng.integers(-4, 9, size=N)))
        busy = rng.integers(0, 400, size=(T, 1, N))
        cases.append(("limits far apart", ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=0,
                                                    max_pods=10**6, min_pods=-(10**6), scale_up_pods=3, scale_down_pods=2),
                      busy, rng.integers(-50, 50, size=N)))
        cases.append(("limits far apart, cool-downs", ksa.Policy(attr_mask=1, scale_up_cool_down=15 * S, scale_down_cool_down=40 * S,
                                                                max_pods=10**6, min_pods=-(10**6), scale_up_pods=7, scale_down_pods=1),
                      busy, rng.integers(-50, 50, size=N)))
        near = rng.integers(0, 400, size=(T, 1, N))
        near[rng.integers(0, 4, size=near.shape) != 0] = 50  # a call every few ticks
        cases.append(("limits a few steps away", ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=0,
                                                           max_pods=12, min_pods=-3), near, rng.integers(-6, 16, size=N)))
        cases.append(("steps that wrap", ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=0, max_pods=2**31 - 1,
                                                   min_pods=-(2**31), scale_up_pods=2**31 - 5, scale_down_pods=2**30 + 3),
                      near, rng.integers(-6, 16, size=N)))
        for name, pol, depth, r0 in cases:
            depth = np.ascontiguousarray(depth, dtype=np.int64)
            st = (r0.astype(np.int32), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
            for chunks in (2, 5, 13):
                for warm in (0, 16):
                    run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"{name} N={N} chunks={chunks} warm={warm}", warm=warm)
            run_and_check(engine, pol, depth, T, 1, N, N, 0, st, 5, f"{name} N={N} with failures", warm=8,
                          fail=random_mask(rng, T, N, 0.15))


def test_saturated_and_idle_queues(engine):
    """The two steady states of a real fleet: a queue that stays above the scale-up threshold (replicas
    at max, ScaleUp called every cool-down period) and one that stays empty (at min, ScaleDown every
    period). The calls of such a stretch are exactly one cool-down apart and their phase is set by its
    first call, so a chunk started from a guess inside the stretch keeps a wrong phase for ever. Summary
    mode works out the true entry states (the stretch's effect on a state has a closed form) and
    replays those chunks a second time; here with cool-downs that do and do not divide the chunk
    length, stretches that begin and end inside chunks, and a random-walk depth (long stretches on one
    side of a threshold)."""
    rng = np.random.default_rng(63)
    T = 900
    for N in (70, 64):
        sat = np.full((T, 1, N), 500)
        idle = np.zeros((T, 1, N), dtype=np.int64)
        mixed = rng.integers(0, 400, size=(T, 1, N))
        mixed[150:520] = 500      # saturated from inside chunk ... to inside chunk ...
        mixed[600:] = 0           # then idle to the end
        per_trace = rng.integers(0, 400, size=(T, 1, N))
        starts = rng.integers(50, 400, size=N)
        for n in range(N):
            per_trace[starts[n]:starts[n] + 350, 0, n] = 300 if n % 2 else 3
        walk = np.clip(np.cumsum(rng.integers(-15, 16, size=(T, 1, N)), axis=0) + 250, 0, 500)
        policies = [ksa.Policy(attr_mask=1),                                                               # cool-downs of 2 and 6 ticks
                    ksa.Policy(attr_mask=1, scale_up_cool_down=35 * S, scale_down_cool_down=55 * S),       # 7 and 11
                    ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=5 * S, max_pods=9),  # 0 and 1
                    ksa.Policy(attr_mask=1, scale_up_cool_down=4000 * S, scale_down_cool_down=13 * S, max_pods=40, min_pods=-7,
                               scale_up_pods=3, scale_down_pods=2),                                        # longer than a chunk
                    ksa.Policy(attr_mask=1, scale_up_messages=10, scale_down_messages=100)]                # thresholds overlap
        for pi, pol in enumerate(policies):
            for name, depth in (("saturated", sat), ("idle", idle), ("mixed", mixed), ("per trace", per_trace), ("walk", walk)):
                depth = np.ascontiguousarray(depth, dtype=np.int64)
                st = (rng.integers(-2, 12, size=N).astype(np.int32), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
                for chunks in (3, 7):
                    run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"{name} policy {pi} N={N} chunks={chunks}", warm=16)
            run_and_check(engine, policies[pi], np.ascontiguousarray(mixed, dtype=np.int64), T, 1, N, N, 0, st, 5,
                          f"mixed policy {pi} N={N} with failures", warm=8, fail=random_mask(rng, T, N, 0.05))
            # the probe that comes before pass 1 of the full replay above 16384 traces, forced on: on the saturated
            # depths it calls the chunked passes off (every trace goes to the plain kernel), on random ones it does not
            for name, depth in (("saturated", sat), ("independent", rng.integers(0, 400, size=(T, 1, N)))):
                took = run_and_check(engine, pol, np.ascontiguousarray(depth, dtype=np.int64), T, 1, N, N, 0, st, 3,
                                     f"{name} policy {pi} N={N} with the probe", probe=1)
                assert took == TP_LAUNCHES + 1


def test_overlapping_thresholds_step_up_and_down_in_one_tick(engine):
    """With scale_up_messages <= scale_down_messages a tick can step up and down: the replica count
    after the tick does not show whether the step up was clamped in between, so a chunk with calls
    must not be shifted by a replica offset (found by tools/fuzz_timepar.py: seed 2, cases 36 and 161)."""
    rng = np.random.default_rng(2)
    T, N = 780, 96
    for pol in (ksa.Policy(attr_mask=1, poll_period=7 * S, scale_up_cool_down=42 * S, scale_down_cool_down=11 * S, scale_up_messages=252,
                           scale_down_messages=366, scale_up_pods=11, scale_down_pods=6, max_pods=5, min_pods=-3),
                ksa.Policy(attr_mask=1, poll_period=7 * S, scale_up_cool_down=35 * S, scale_down_cool_down=0, scale_up_messages=10,
                           scale_down_messages=386, scale_up_pods=3, scale_down_pods=1, max_pods=6, min_pods=-3)):
        depth = np.ascontiguousarray(rng.integers(0, 800, size=(T, 1, N)), dtype=np.int64)
        st = (rng.integers(-3, 12, size=N).astype(np.int32), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
        for chunks in (2, 5, 30):
            run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"overlapping thresholds chunks={chunks} {pol}", warm=0)


def test_random_shapes_policies_and_depths(engine):
    """A short run of tools/fuzz_timepar.py: random shapes, policies (adversarial and ordinary), depths
    mixed per trace (independent, random walks, saturated / idle / quiet stretches), masks, chunk
    counts, warm-ups and probe settings through the full and the summary entry point."""
    from tools import fuzz_timepar

    assert fuzz_timepar.run(120, 11, engine) == 0
    assert fuzz_timepar.run(120, 12, engine) == 0


def test_two_streams_share_the_scratch_safely(engine):
    """The chunked passes keep their records in scratch arrays inside the library: one such call per
    GPU at a time. A call that finds the scratch in use by a call on another stream takes the plain
    kernels instead; calls on one stream are ordered anyway. Both give the oracle's results."""
    import torch

    gen_a, gen_b = ksa.GenParams(seed=101), ksa.GenParams(seed=202)
    N, T = 4096, 8192
    pol = ksa.Policy(attr_mask=1)
    cases = []
    for gen in (gen_a, gen_b):
        depth = oracle_lib.gen_traces(gen.to_c(), 0, T, 1, 0, N, N)
        st = oracle_lib.fresh_state(gen.seed, 0, N, 1, 5)
        ref = oracle_lib.replay(pol.to_c(), depth, T, 1, N, N, 0, *st, want_summary=True)
        cases.append((torch.from_numpy(depth).to(engine.dev), st, ref))
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(device=engine.dev), torch.cuda.Stream(device=engine.dev)]
    for order in ((0, 1), (1, 1), (0, 0)):  # two streams, then the same stream twice
        outs = []
        for (d, st, ref), si in zip(cases, order):
            with torch.cuda.stream(streams[si]):
                s1, s2 = engine.state_from_host(*st), engine.state_from_host(*st)
                r, e = engine.replay_full(pol, d, T, 1, N, N, 0, s1)
                summ = engine.replay_summary(pol, d, T, 1, N, N, 0, s2)
                outs.append((r, e, summ, s1, s2, ref))
        torch.cuda.synchronize()
        for r, e, summ, s1, s2, ref in outs:
            assert np.array_equal(r.cpu().numpy(), ref.replicas) and np.array_equal(e.cpu().numpy(), ref.events)
            assert summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.summary.tobytes()
            for s in (s1, s2):
                assert np.array_equal(s.replicas.cpu().numpy(), ref.state_replicas)
                assert np.array_equal(s.last_up.cpu().numpy(), ref.state_last_up)
                assert np.array_equal(s.last_down.cpu().numpy(), ref.state_last_down)


def test_limits_at_the_ends_of_int32(engine):
    """The reference adds a step in int32 and wraps before it clamps (scale/scale.go:66, 94). With
    max_pods = 2^31 - 1 a step from just below the limit passes INT_MAX and comes out negative; the
    closed forms for saturated / idle stretches and the shift of a chunk add in 64 bits and must stay
    away from such policies (found by tools/fuzz_timepar.py: seed 82, case 76)."""
    rng = np.random.default_rng(82)
    T, N = 975, 96
    pol = ksa.Policy(attr_mask=1, scale_up_cool_down=20 * S, scale_down_cool_down=15 * S, scale_up_messages=183,
                     scale_down_messages=20, scale_up_pods=8, scale_down_pods=5, max_pods=2**31 - 1, min_pods=-(2**31))
    r0 = np.where(rng.integers(0, 2, size=N) == 0, 2**31 - 1 - rng.integers(0, 40, size=N),
                  -(2**31) + rng.integers(0, 40, size=N)).astype(np.int32)
    st = (r0, np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
    sat = np.full((T, 1, N), 500, dtype=np.int64)
    idle = np.zeros((T, 1, N), dtype=np.int64)
    mixed = np.ascontiguousarray(rng.integers(0, 400, size=(T, 1, N)), dtype=np.int64)
    mixed[200:700] = 500
    for name, depth in (("saturated", sat), ("idle", idle), ("mixed", mixed)):
        for chunks in (3, 12):
            run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"limits at the ends of int32, {name}, chunks={chunks}", warm=64)


def test_replica_offsets_wider_than_int32(engine):
    """A chunk replayed from a guess whose replica count is off by 2^31 or more (the trace starts near
    INT_MIN, a negative scale_up_pods wraps it past the limit in the first ticks and the clamp brings it
    to max_pods; the chunks after that see no call and were guessed from the caller's count). Such a
    chunk is still right after a shift, and the shift has to be made in 64 bits all the way: the sum of
    replicas is the one field that is not taken mod 2^32 (found by tools/fuzz_timepar.py: seed 92)."""
    rng = np.random.default_rng(92)
    T, N = 1282, 96
    pol = ksa.Policy(attr_mask=1, scale_up_cool_down=0, scale_down_cool_down=13 * S, scale_up_messages=109,
                     scale_down_messages=43, scale_up_pods=-3, scale_down_pods=11, max_pods=6, min_pods=0)
    r0 = np.where(rng.integers(0, 2, size=N) == 0, -(2**31) + rng.integers(0, 3, size=N), rng.integers(0, 7, size=N)).astype(np.int32)
    st = (r0, np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
    depth = np.ascontiguousarray(rng.integers(44, 109, size=(T, 1, N)), dtype=np.int64)  # between the thresholds: no call
    depth[:5] = 200
    for chunks in (4, 30):
        for warm in (0, 24):
            run_and_check(engine, pol, depth, T, 1, N, N, 0, st, chunks, f"offsets wider than int32, chunks={chunks}, warm={warm}", warm=warm)
