"""Out-of-bounds writes, checked with canaries (compute-sanitizer is closed on this GPU pool): every
output array is embedded in a larger buffer filled with a marker; after the call the marker must be
intact before the array, after its last row, and in the padding columns of every row. Shapes walk
through the edges of the tick loop (T around the ring depth and its multiples) and of the trace
axis (N around the traces-per-thread of each kernel variant)."""
from __future__ import annotations

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from tests import oracle_lib
from tests.test_gpu_parity import full_variants

pytestmark = pytest.mark.gpu
S = 10**9


@pytest.mark.parametrize("A", [1, 3])
def test_replay_never_writes_outside_its_outputs(engine, A):
    import torch

    rng = np.random.default_rng(A)
    pol = ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=120, scale_down_messages=30,
                     scale_up_cool_down=10 * S, scale_down_cool_down=15 * S, max_pods=9)
    guard = 64  # rows of marker before and after
    try:
        for T in (1, 2, 3, 5, 9, 17):
            for N, ld in ((1, 16), (7, 16), (9, 16), (17, 32), (33, 48), (129, 144)):
                depth_np = rng.integers(0, 200, size=(T, A, ld), dtype=np.int64)
                st = oracle_lib.fresh_state(5, 0, N, 1, 9)
                ref = oracle_lib.replay(pol.to_c(), depth_np, T, A, N, ld, 0, *st, want_summary=True)
                depth = torch.from_numpy(depth_np).to(engine.dev)
                for fid in full_variants(engine):
                    engine.set_variant(fid, 0)
                    rbuf = torch.full((T + 2 * guard, ld), -7777, dtype=torch.int32, device=engine.dev)
                    ebuf = torch.full((T + 2 * guard, ld), 0xEE, dtype=torch.uint8, device=engine.dev)
                    sbuf = torch.full((N + 2 * guard, 64), 0xEE, dtype=torch.uint8, device=engine.dev)
                    s1, s2 = engine.state_from_host(*st), engine.state_from_host(*st)
                    engine.replay_full(pol, depth, T, A, N, ld, 0, s1, replicas=rbuf[guard:guard + T],
                                       events=ebuf[guard:guard + T])
                    summ = sbuf[guard:guard + N]
                    engine.lib.ksa_summary_clear(summ.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
                    engine.replay_summary(pol, depth, T, A, N, ld, 0, s2, summary=summ)
                    label = f"A={A} T={T} N={N} ld={ld} variant={fid}"
                    assert (rbuf[:guard] == -7777).all() and (rbuf[guard + T:] == -7777).all(), label
                    assert (ebuf[:guard] == 0xEE).all() and (ebuf[guard + T:] == 0xEE).all(), label
                    assert (rbuf[guard:guard + T, N:] == -7777).all() and (ebuf[guard:guard + T, N:] == 0xEE).all(), label
                    assert (sbuf[:guard] == 0xEE).all() and (sbuf[guard + N:] == 0xEE).all(), label
                    assert np.array_equal(rbuf[guard:guard + T, :N].cpu().numpy(), ref.replicas[:, :N]), label
                    assert np.array_equal(ebuf[guard:guard + T, :N].cpu().numpy(), ref.events[:, :N]), label
                    assert summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.summary.tobytes(), label
    finally:
        engine.set_variant(0, 0)


def test_ingest_and_generators_never_write_outside(engine):
    import torch

    g = ksa.GenParams(seed=3, low_max=25, high_min=250, high_max=1000)
    for T, N, ld, A in ((1, 1, 16, 1), (65, 33, 48, 3), (257, 31, 32, 2), (300, 70, 80, 1)):
        buf = torch.full((T + 8, A, ld), -5, dtype=torch.int64, device=engine.dev)
        engine.gen_traces(g, 0, T, A, 0, N, ld, out=buf[4:4 + T])
        assert (buf[:4] == -5).all() and (buf[4 + T:] == -5).all() and (buf[4:4 + T, :, N:] == -5).all()
        fb = torch.full((T + 8, ld), 0xEE, dtype=torch.uint8, device=engine.dev)
        engine.gen_faults(3, 2, 2, 2, 0, T, 0, N, ld, out=fb[4:4 + T])
        assert (fb[:4] == 0xEE).all() and (fb[4 + T:] == 0xEE).all() and (fb[4:4 + T, N:] == 0xEE).all()
        src = torch.randint(-9, 9, (N, T + 3), dtype=torch.int64, device=engine.dev)
        buf.fill_(-5)
        engine.ingest_transpose(src, T, N, A, A - 1, ld, out=buf[4:4 + T])
        assert (buf[:4] == -5).all() and (buf[4 + T:] == -5).all() and (buf[4:4 + T, :, N:] == -5).all()
        if A > 1:
            assert (buf[4:4 + T, : A - 1] == -5).all()
        assert torch.equal(buf[4:4 + T, A - 1, :N], src[:, :T].t())


def test_summary_variants_sweep_and_reduce_never_write_outside(engine, monkeypatch):
    """Canaries around the summary records for every summary variant (register ring, ordered step,
    cp.async.bulk staging), around the [P][N] output of both sweep kernels, and around the three
    outputs of ksa_sweep_reduce; shapes with ragged tails and a last policy tile that is not full."""
    import torch

    from tests.helpers import variants_present
    from tests.test_gpu_parity import summary_variants

    rng = np.random.default_rng(5)
    guard = 16
    pol = ksa.Policy(attr_mask=1, scale_up_messages=120, scale_down_messages=30, scale_up_cool_down=10 * S,
                     scale_down_cool_down=15 * S, max_pods=9)
    try:
        for T, N, ld in ((1, 1, 16), (9, 130, 144), (40, 257, 272), (33, 128, 128)):
            depth_np = rng.integers(0, 200, size=(T, 1, ld), dtype=np.int64)
            depth = torch.from_numpy(depth_np).to(engine.dev)
            st = oracle_lib.fresh_state(5, 0, N, 1, 9)
            ref = oracle_lib.replay(pol.to_c(), depth_np, T, 1, N, ld, 0, *st, want_full=False, want_summary=True)
            for sv in summary_variants(engine):
                engine.set_variant(0, sv)
                sbuf = torch.full((N + 2 * guard, 64), 0xEE, dtype=torch.uint8, device=engine.dev)
                summ = sbuf[guard:guard + N]
                engine.lib.ksa_summary_clear(summ.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
                engine.replay_summary(pol, depth, T, 1, N, ld, 0, engine.state_from_host(*st), summary=summ)
                label = f"T={T} N={N} ld={ld} summary variant {sv}"
                assert (sbuf[:guard] == 0xEE).all() and (sbuf[guard + N:] == 0xEE).all(), label
                assert summ.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.summary.tobytes(), label
            engine.set_variant(0, 0)
            # sweep: 7 policies (tile of 4 + ragged 3), mixed narrow / general tiles
            pols = [ksa.Policy(attr_mask=1, scale_up_messages=60 + 20 * i, scale_down_messages=(70 if i == 5 else 10 + i),
                               scale_up_cool_down=(i % 3) * 5 * S, max_pods=4 + i) for i in range(7)]
            P = len(pols)
            r0 = torch.from_numpy(st[0]).to(engine.dev)
            want = [oracle_lib.replay(p.to_c(), depth_np, T, 1, N, ld, 0, *st, want_full=False, want_summary=True).summary
                    for p in pols]
            for general in ("0", "1"):
                monkeypatch.setenv("KSA_SWEEP_GENERAL", general)
                for sv in variants_present(engine, 4, (2, 9, 16, 26, 27, 33, 37)):
                    monkeypatch.setenv("KSA_SWEEP_VARIANT", str(sv))
                    big = torch.full((P * N + 2 * guard, 64), 0xEE, dtype=torch.uint8, device=engine.dev)
                    out = big[guard:guard + P * N].view(P, N, 64)
                    engine.replay_sweep(pols, depth, T, 1, N, ld, r0, summary=out)
                    label = f"T={T} N={N} sweep variant {sv} general={general}"
                    assert (big[:guard] == 0xEE).all() and (big[guard + P * N:] == 0xEE).all(), label
                    got = out.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(P, N)
                    assert all(got[i].tobytes() == want[i].tobytes() for i in range(P)), label
            monkeypatch.delenv("KSA_SWEEP_GENERAL")
            monkeypatch.delenv("KSA_SWEEP_VARIANT")
            # reduce
            pc = torch.full((P + 2 * guard,), -77, dtype=torch.int64, device=engine.dev)
            bp = torch.full((N + 2 * guard,), -77, dtype=torch.int32, device=engine.dev)
            bc = torch.full((N + 2 * guard,), -77, dtype=torch.int64, device=engine.dev)
            pc[guard:guard + P] = 0
            w = ksa._abi.KsaCost(1, 2, 3, 4)
            import ctypes as C
            ksa._abi.check(engine.lib.ksa_sweep_reduce(out.data_ptr(), P, N, C.byref(w), pc[guard:].data_ptr(),
                                                       bp[guard:].data_ptr(), bc[guard:].data_ptr(),
                                                       torch.cuda.current_stream().cuda_stream))
            for buf, n in ((pc, P), (bp, N), (bc, N)):
                assert (buf[:guard] == -77).all() and (buf[guard + n:] == -77).all(), f"reduce T={T} N={N}"
            assert (bp[guard:guard + N] >= 0).all() and (bp[guard:guard + N] < P).all()
    finally:
        engine.set_variant(0, 0)
