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

pytestmark = pytest.mark.gpu
S = 10**9
FULL_IDS = [-2, 1, 14, 16, 19, 20, 21, 23, 27, 30, 31, 32, 34, 36, 38, 41]


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
                for fid in FULL_IDS:
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
