#!/usr/bin/env python
"""A small pass through every kernel of the library, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck python tests/sanitize_case.py

Shapes are chosen to hit the edges: ragged N (scalar tail launch), T not a multiple of the ring
depth, T smaller than the ring, padded pitch, every A, every kernel variant, both sweep layouts,
the failure-mask kernels, the ingest kernels and the host pipeline. Results are compared with the
oracle so that the run also fails on wrong answers, not only on bad addresses.
"""
from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import kube_sqs_autoscaler_b200 as ksa  # noqa: E402
from tests import oracle_lib  # noqa: E402

S = 10**9


def main():
    eng = ksa.ReplayEngine(0)
    rng = np.random.default_rng(1)
    checked = 0
    for A in (1, 2, 3):
        pol = ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=120, scale_down_messages=30,
                         scale_up_cool_down=12 * S, scale_down_cool_down=20 * S, max_pods=9)
        for T, N, ld in ((1, 1, 1), (3, 37, 48), (17, 515, 528), (9, 64, 64)):
            depth = rng.integers(0, 200, size=(T, A, ld), dtype=np.int64)
            fail = rng.choice(np.array([0, 0, 0, 1, 2, 4, 7], dtype=np.uint8), size=(T, ld))
            st = oracle_lib.fresh_state(5, 0, N, 1, 9)
            ref = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, 0, *st, want_summary=True)
            reff = oracle_lib.replay(pol.to_c(), depth, T, A, N, ld, 0, *st, want_summary=True, fail=fail)
            d = torch.from_numpy(depth).to(eng.dev)
            f = torch.from_numpy(fail).to(eng.dev)
            full_ids = [-2] + [i for i in range(1, 64) if eng.variant_name(0, i)]
            sum_ids = [-2] + [i for i in range(1, 64) if eng.variant_name(1, i)]
            for k, fid in enumerate(full_ids):
                eng.set_variant(fid, sum_ids[k % len(sum_ids)])
                s1, s2 = eng.state_from_host(*st), eng.state_from_host(*st)
                r, e = eng.replay_full(pol, d, T, A, N, ld, 0, s1)
                sm = eng.replay_summary(pol, d, T, A, N, ld, 0, s2)
                assert np.array_equal(r.cpu().numpy()[:, :N], ref.replicas[:, :N])
                assert np.array_equal(e.cpu().numpy()[:, :N], ref.events[:, :N])
                assert sm.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == ref.summary.tobytes()
                checked += 1
            eng.set_variant(0, 0)
            for fv in ["0"] + [str(i) for i in range(1, 32) if eng.variant_name(2, i) or eng.variant_name(3, i)]:
                os.environ["KSA_FAULTS_VARIANT"] = fv
                s1, s2 = eng.state_from_host(*st), eng.state_from_host(*st)
                r, e = eng.replay_full(pol, d, T, A, N, ld, 0, s1, fail=f)
                sm = eng.replay_summary(pol, d, T, A, N, ld, 0, s2, fail=f)
                assert np.array_equal(r.cpu().numpy()[:, :N], reff.replicas[:, :N])
                assert np.array_equal(e.cpu().numpy()[:, :N], reff.events[:, :N])
                assert sm.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1).tobytes() == reff.summary.tobytes()
                checked += 1
            os.environ.pop("KSA_FAULTS_VARIANT", None)
            h = eng.traj_hash(r, e, T, N, ld)
            assert np.array_equal(h.cpu().numpy().view(np.uint64), reff.summary["hash"])
            # sweep, both layouts and the built-in table
            pols = [ksa.Policy(attr_mask=(1 << A) - 1, scale_up_messages=60 + 20 * i, scale_down_messages=10 + 5 * i,
                               scale_up_cool_down=(i % 3) * 5 * S, scale_down_cool_down=(i % 4) * 5 * S,
                               max_pods=4 + i) for i in range(7)]
            r0 = torch.from_numpy(st[0]).to(eng.dev)
            want = [oracle_lib.replay(p.to_c(), depth, T, A, N, ld, 0, *st, want_full=False, want_summary=True).summary
                    for p in pols]
            for sv in [str(i) for i in (0, 2, 5, 8, 9, 16, 20, 26, 33, 37) if eng.variant_name(4, i)]:
                os.environ["KSA_SWEEP_VARIANT"] = sv
                got = eng.replay_sweep(pols, d, T, A, N, ld, r0).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(len(pols), N)
                for i in range(len(pols)):
                    assert got[i].tobytes() == want[i].tobytes(), (sv, i)
                checked += 1
            os.environ.pop("KSA_SWEEP_VARIANT", None)
            out = torch.empty((len(pols), N, 64), dtype=torch.uint8, device=eng.dev)
            eng.replay_sweep_static(pols, d, T, A, N, ld, r0, out)
            assert out.cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(len(pols), N)[3].tobytes() == want[3].tobytes()
    # generators
    for fam in (ksa.GEN_BANDS, ksa.GEN_WALK, ksa.GEN_ADVERSARIAL):
        g = ksa.GenParams(seed=3, family=fam, low_max=25, high_min=250, high_max=1000, overflow_log2=4)
        w = eng.gen_walk_init(g, 7, 131) if fam == ksa.GEN_WALK else None
        hw = oracle_lib.gen_walk_init(g.to_c(), 7, 131) if fam == ksa.GEN_WALK else None
        got = eng.gen_traces(g, 5, 67, 3, 7, 131, 144, walk=w).cpu().numpy()
        assert np.array_equal(got[:, :, :131], oracle_lib.gen_traces(g.to_c(), 5, 67, 3, 7, 131, 144, walk=hw)[:, :, :131])
    assert np.array_equal(eng.gen_faults(3, 2, 3, 1, 5, 67, 7, 131, 144).cpu().numpy()[:, :131],
                          oracle_lib.gen_faults(3, 2, 3, 1, 5, 67, 7, 131, 144)[:, :131])
    # ingest
    for dt in (np.int64, np.int32):
        src = rng.integers(-1000, 1000, size=(70, 301), dtype=dt)
        for shape in range(12):
            os.environ["KSA_TRANSPOSE_SHAPE"] = str(shape)
            try:
                got = eng.ingest_transpose(torch.from_numpy(src).to(eng.dev), 300, 70, 2, 1, 80).cpu().numpy()
            except ksa.KsaError:
                continue  # a tile shape only the tuning build has
            assert np.array_equal(got[:, 1, :70], src[:, :300].T.astype(np.int64))
        os.environ.pop("KSA_TRANSPOSE_SHAPE", None)
    texts = [b"", b"12", b"-7", b"+", b"9223372036854775807", b"9223372036854775808", b"1 2", b"00000001", b"123456789"]
    for width in (16, 32, 21):
        rec = np.zeros((len(texts), width), dtype=np.uint8)
        for i, t in enumerate(texts):
            rec[i, : min(len(t), width)] = np.frombuffer(t[:width], dtype=np.uint8)
        v, fl = eng.ingest_decimal(torch.from_numpy(rec).to(eng.dev))
        hv, hf = oracle_lib.atoi_records(rec)
        assert np.array_equal(v.cpu().numpy(), hv) and np.array_equal(fl.cpu().numpy(), hf)
    # host pipeline
    pol = ksa.Policy(attr_mask=1)
    T, N = 23, 301
    depth = ksa.pinned_empty((T, 1, N), np.int64)
    depth[...] = rng.integers(0, 300, size=(T, 1, N))
    st = oracle_lib.fresh_state(2, 0, N, 1, 5)
    ref = oracle_lib.replay(pol.to_c(), depth, T, 1, N, N, 0, *st, want_summary=True)
    pipe = ksa.HostPipeline(0, N, 1, 5)
    rr, ee = ksa.pinned_empty((T, N), np.int32), ksa.pinned_empty((T, N), np.uint8)
    summ = np.empty(N, dtype=ksa.SUMMARY_DTYPE)
    a, b, c = (np.array(x, copy=True) for x in st)
    pipe.replay(pol, depth, 0, a, b, c, rr, ee, summ)
    pipe.close()
    assert np.array_equal(rr, ref.replicas) and np.array_equal(ee, ref.events) and summ.tobytes() == ref.summary.tobytes()
    torch.cuda.synchronize()
    print(f"sanitize_case ok: {checked} replay/sweep configurations + generators + ingest + pipeline")


if __name__ == "__main__":
    main()
