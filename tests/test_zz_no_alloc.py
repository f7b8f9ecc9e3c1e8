"""ksa_replay_sweep_static — the policy sweep for callers that cannot allocate.

Some hosts of this library have no safe way to obtain device memory at call time: the GPU's memory
already belongs to another framework's pool, the host is Go over cgo with no CUDA allocator of its
own, or the sweep runs in a loop that a cudaMalloc / cudaFree would stall. For them the library
keeps the sweep's kernel-side policy table in a fixed array of its own (4096 policies). This file
checks, through the C ABI, that the entry point returns exactly what ksa_replay_sweep returns, on
two calls in a row that share that array (the second with fewer and different policies, so stale
records from the first call would show), that it leaves the amount of free device memory
unchanged, and that it refuses a grid larger than its table.

Named test_zz_* so that pytest collects it after every other test file; it also passes when run on
its own in a fresh process.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from kube_sqs_autoscaler_b200 import _abi
from kube_sqs_autoscaler_b200.engine import policy_array

pytestmark = pytest.mark.gpu
S = 10**9


def _grid(count: int, shift: int) -> list:
    """`count` distinct policies; `shift` moves the thresholds so that two grids share no record."""
    out = []
    for i in range(count):
        out.append(ksa.Policy(attr_mask=1, scale_up_messages=60 + 17 * ((i + shift) % 19),
                              scale_down_messages=5 + 9 * ((i * 7 + shift) % 13),
                              scale_up_cool_down=((i + shift) % 5) * 5 * S,
                              scale_down_cool_down=((i * 3 + shift) % 7) * 5 * S,
                              min_pods=1 + (i + shift) % 3, max_pods=4 + (i * 5 + shift) % 40,
                              scale_up_pods=1 + (i + shift) % 4, scale_down_pods=1 + (i * 2 + shift) % 3))
    return out


def test_static_sweep_matches_sweep_on_two_calls_in_a_row(engine):
    import torch

    lib = engine.lib
    gen = ksa.GenParams(seed=0xA110C, family=ksa.GEN_WALK, high_max=500, walk_step=15)
    T, N = 400, 1000
    walk = engine.gen_walk_init(gen, 0, N)
    depth = engine.gen_traces(gen, 0, T, 1, 0, N, N, walk=walk)
    r0 = engine.new_state(N, seed=gen.seed, lo=1, hi=5).replicas
    stream = torch.cuda.current_stream().cuda_stream

    first, second = _grid(61, 0), _grid(9, 1000)  # second call: fewer, different policies
    # every buffer either entry point touches is created here, before the calls under test
    bufs = {}
    for name, pols in (("first", first), ("second", second)):
        P = len(pols)
        bufs[name] = (
            policy_array(pols), P,
            torch.full((P, N, 64), 0xAB, dtype=torch.uint8, device=engine.dev),   # static entry point
            torch.full((P, N, 64), 0xCD, dtype=torch.uint8, device=engine.dev),   # existing entry point
            torch.empty(P * _abi.POLICY_K_BYTES, dtype=torch.uint8, device=engine.dev),
        )
    torch.cuda.synchronize()

    for call, name in enumerate(("first", "second")):
        arr, P, out_static, out_ref, table = bufs[name]
        free_before, _ = torch.cuda.mem_get_info()
        _abi.check(lib.ksa_replay_sweep_static(arr, P, depth.data_ptr(), T, 1, N, N, r0.data_ptr(),
                                               out_static.data_ptr(), stream))
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        if call > 0:  # the very first launch of a kernel may make the driver load its code
            assert free_after == free_before, "the static sweep changed the amount of free device memory"
        _abi.check(lib.ksa_replay_sweep(arr, P, table.data_ptr(), depth.data_ptr(), T, 1, N, N,
                                        r0.data_ptr(), out_ref.data_ptr(), stream))
        torch.cuda.synchronize()
        # integer work: the tolerance is zero, byte for byte
        assert torch.equal(out_static, out_ref), f"call {call + 1}: static sweep differs from ksa_replay_sweep"
        got = out_static.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
        assert got["up_applied"].sum() > 0 and got["down_applied"].sum() > 0 and got["up_cool"].sum() > 0
        if name == "second":
            # and directly against the CPU oracle, policy by policy (the sibling entry point is
            # oracle-tested on other grids; this makes the check of the static one first-hand)
            from tests import oracle_lib

            d_host = depth.cpu().numpy()
            st = (r0.cpu().numpy(), np.zeros(N, dtype=np.int64), np.zeros(N, dtype=np.int64))
            pols = second
            for i, pol in enumerate(pols):
                ref = oracle_lib.replay(pol.to_c(), d_host, T, 1, N, N, 0, *st, want_full=False, want_summary=True)
                assert got[i].tobytes() == ref.summary.tobytes(), f"static sweep, policy {i}: differs from the oracle"


def test_static_sweep_rejects_a_grid_larger_than_its_table(engine):
    import torch

    lib = engine.lib
    T, N = 4, 16
    depth = torch.zeros((T, 1, N), dtype=torch.int64, device=engine.dev)
    r0 = torch.ones(N, dtype=torch.int32, device=engine.dev)
    P = _abi.SWEEP_STATIC_MAX_POLICIES + 1
    arr = (_abi.KsaPolicy * P)(*([ksa.Policy(attr_mask=1).to_c()] * P))
    out = torch.zeros((1, N, 64), dtype=torch.uint8, device=engine.dev)  # never written
    rc = lib.ksa_replay_sweep_static(arr, P, depth.data_ptr(), T, 1, N, N, r0.data_ptr(), out.data_ptr(), None)
    assert rc == _abi.KSA_ELIMIT
    msg = lib.ksa_last_error().decode()
    assert str(_abi.SWEEP_STATIC_MAX_POLICIES) in msg and "ksa_replay_sweep" in msg
    torch.cuda.synchronize()
    assert int(out.sum()) == 0
    # at the limit it still works
    P = _abi.SWEEP_STATIC_MAX_POLICIES
    out = torch.zeros((P, N, 64), dtype=torch.uint8, device=engine.dev)
    _abi.check(lib.ksa_replay_sweep_static(arr, P, depth.data_ptr(), T, 1, N, N, r0.data_ptr(), out.data_ptr(), None))
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
    # depth 0 under the default policy: every tick is a scale-down still inside the 6-tick cool-down
    assert (got["final_replicas"] == 1).all() and (got["down_cool"] == T).all()
