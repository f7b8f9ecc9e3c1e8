"""Ingest: trace-major -> tick-major transposition and the strconv.Atoi decode
of SQS attribute strings (sqs/sqs.go:58). CPU tests pin the C oracle to the golden Atoi vectors and
to the independent Python restatement; GPU tests compare the CUDA kernels with the oracle / numpy,
bit for bit, through the C ABI."""
from __future__ import annotations

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import kube_sqs_autoscaler_b200 as ksa
from oracle import policy_py
from tests import oracle_lib
from tests.helpers import load_golden

ATOI = load_golden("atoi_vectors.json")


def records(texts: list[bytes], width: int) -> np.ndarray:
    out = np.zeros((len(texts), width), dtype=np.uint8)
    for i, t in enumerate(texts):
        t = t[:width]
        out[i, : len(t)] = np.frombuffer(t, dtype=np.uint8)
    return out


@pytest.mark.parametrize("group", ["reference", "derived"])
def test_atoi_golden_vectors(group):
    cases = ATOI[group]
    texts = [c["text"].encode("utf-8") for c in cases]
    values, fail = oracle_lib.atoi_records(records(texts, 32))
    for c, t, v, f in zip(cases, texts, values, fail):
        assert policy_py.atoi(t) == c["value"], c
        if c["value"] is None:
            assert f == 1 and v == 0, c
        else:
            assert f == 0 and int(v) == c["value"], c


ALPHABET = "0123456789+- _aex."


@settings(max_examples=400, deadline=None)
@given(st.lists(st.one_of(st.text(alphabet=ALPHABET, max_size=22),
                          st.integers(-(2**64), 2**64).map(str)), min_size=1, max_size=20),
       st.sampled_from([8, 16, 20, 32]))
def test_atoi_c_oracle_equals_python_restatement(texts, width):
    raw = [t.encode("ascii")[:width] for t in texts]
    values, fail = oracle_lib.atoi_records(records(raw, width))
    for t, v, f in zip(raw, values, fail):
        want = policy_py.atoi(t)
        assert (want is None) == (f == 1), t
        assert int(v) == (want or 0), t


def test_reference_sum_through_the_decoder():
    """sqs/sqs_test.go:19-25 end to end: "10","10","10" -> Atoi -> sum 30."""
    v, f = oracle_lib.atoi_records(records([b"10", b"10", b"10"], 16))
    assert f.sum() == 0 and oracle_lib.num_messages(v) == 30


# ---- GPU ------------------------------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("width", [16, 32, 20, 7])
def test_decimal_kernel_matches_oracle(engine, width):
    import torch

    rng = np.random.default_rng(width)
    texts = [c["text"].encode("utf-8") for c in ATOI["reference"] + ATOI["derived"]]
    for _ in range(20000):
        kind = rng.integers(0, 6)
        if kind < 3:
            texts.append(str(int(rng.integers(0, 10 ** int(rng.integers(1, 7))))).encode())
        elif kind == 3:
            texts.append(str(int(rng.integers(-(2**63), 2**63 - 1, endpoint=True))).encode())
        elif kind == 4:
            texts.append(bytes(rng.choice(np.frombuffer(ALPHABET.encode(), dtype=np.uint8), size=int(rng.integers(0, 12)))))
        else:
            texts.append(str(int(rng.integers(2**63 - 5, 2**63 + 5, dtype=np.uint64))).encode())
    rec = records(texts, width)
    want_v, want_f = oracle_lib.atoi_records(rec)
    got_v, got_f = engine.ingest_decimal(torch.from_numpy(rec).to(engine.dev))
    assert np.array_equal(got_v.cpu().numpy(), want_v)
    assert np.array_equal(got_f.cpu().numpy(), want_f)
    assert 0 < want_f.sum() < len(texts)
    # unaligned base pointer -> byte-load path, same results
    big = torch.zeros(rec.size + 1, dtype=torch.uint8, device=engine.dev)
    big[1:] = torch.from_numpy(rec.reshape(-1)).to(engine.dev)
    got_v2, got_f2 = engine.ingest_decimal(big[1:].view(len(texts), width))
    assert np.array_equal(got_v2.cpu().numpy(), want_v) and np.array_equal(got_f2.cpu().numpy(), want_f)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int64, np.int32])
@pytest.mark.parametrize("T,N,A,a", [(64, 64, 1, 0), (1, 1, 1, 0), (130, 67, 3, 1), (1000, 257, 2, 1), (63, 4100, 1, 0)])
def test_transpose_kernel_matches_numpy(engine, dtype, T, N, A, a):
    import torch

    rng = np.random.default_rng(T * 7 + N)
    src_ld = T + (3 if T % 2 else 0)
    lo, hi = (-(2**63), 2**63 - 1) if dtype == np.int64 else (-(2**31), 2**31 - 1)
    src = rng.integers(lo, hi, size=(N, src_ld), dtype=dtype, endpoint=True)
    ld = (N + 15) // 16 * 16
    out = torch.full((T, A, ld), -12345, dtype=torch.int64, device=engine.dev)
    engine.ingest_transpose(torch.from_numpy(src).to(engine.dev), T, N, A, a, ld, out=out)
    got = out.cpu().numpy()
    assert np.array_equal(got[:, a, :N], src[:, :T].T.astype(np.int64))
    mask = np.ones((T, A, ld), dtype=bool)
    mask[:, a, :N] = False
    assert (got[mask] == -12345).all(), "wrote outside its plane / its N columns"


@pytest.mark.gpu
def test_ingest_takes_every_trace_count_the_replay_takes(engine):
    """The transposes put the trace axis on gridDim.y (65535 tiles); more traces than one launch
    holds go in slabs. One more than a slab of the int32 tile (64 ticks x 128 traces), of the int64
    tile (256 x 32) and of the mask tile (64 x 64), a few ticks each: equal to numpy."""
    import torch

    T = 3
    for dtype, N in ((np.int32, 65535 * 128 + 77), (np.int64, 65535 * 32 + 5)):
        src = (np.arange(N * T, dtype=np.int64) * 2654435761 % 100003 - 50000).astype(dtype).reshape(N, T)
        got = engine.ingest_transpose(torch.from_numpy(src).to(engine.dev), T, N, 1, 0, N)
        want = torch.from_numpy(src).to(engine.dev).t().to(torch.int64)
        assert bool(torch.equal(got[:, 0, :], want)), dtype
        del got, want
    N = 65535 * 64 + 9
    m = (np.arange(N * T, dtype=np.int64) * 40503 % 251 % 8).astype(np.uint8).reshape(N, T)
    got = engine.ingest_transpose_mask(torch.from_numpy(m).to(engine.dev), T, N, N)
    assert bool(torch.equal(got, torch.from_numpy(m).to(engine.dev).t()))


@pytest.mark.gpu
def test_recorded_strings_to_trajectory(engine):
    """The whole ingest chain on a small recording: per-queue histories of attribute STRINGS (what
    SQS returns, sqs/sqs.go:58) -> Atoi decode -> transpose to tick-major -> replay with the parse
    failures as failed polls; equals the oracle fed the same strings one tick at a time."""
    import torch

    rng = np.random.default_rng(99)
    T, N = 200, 150
    vals = rng.integers(0, 400, size=(N, T))
    texts = [str(int(v)).encode() for v in vals.reshape(-1)]
    bad = rng.choice(N * T, size=300, replace=False)
    for i in bad:
        texts[i] = [b"", b"n/a", b"12 ", b"1_0", b"9223372036854775808"][int(rng.integers(0, 5))]
    rec = records(texts, 32)
    values, fail = engine.ingest_decimal(torch.from_numpy(rec).to(engine.dev))
    ld = 160
    depth = engine.ingest_transpose(values.view(N, T), T, N, 1, 0, ld)
    # the failure bytes are trace-major too
    fail_t = engine.ingest_transpose_mask(fail.view(N, T), T, N, ld)
    assert torch.equal(fail_t[:, :N], fail.view(N, T).t())
    pol = ksa.Policy(attr_mask=1)
    state = engine.new_state(N, seed=3, lo=1, hi=5)
    replicas, events = engine.replay_full(pol, depth, T, 1, N, ld, 0, state, fail=fail_t)
    # oracle: decode on the host, per tick
    hv, hf = oracle_lib.atoi_records(rec)
    d_host = np.zeros((T, 1, ld), dtype=np.int64)
    d_host[:, 0, :N] = hv.reshape(N, T).T
    f_host = np.zeros((T, ld), dtype=np.uint8)
    f_host[:, :N] = hf.reshape(N, T).T
    ref = oracle_lib.replay(pol.to_c(), d_host, T, 1, N, ld, 0, *oracle_lib.fresh_state(3, 0, N, 1, 5), fail=f_host)
    assert np.array_equal(replicas.cpu().numpy()[:, :N], ref.replicas[:, :N])
    assert np.array_equal(events.cpu().numpy()[:, :N], ref.events[:, :N])
    assert (ref.events[:, :N] == 0xC0).sum() == 300


@pytest.mark.gpu
@pytest.mark.parametrize("T,N,ld", [(1, 1, 16), (63, 65, 80), (200, 150, 160), (64, 64, 64)])
def test_mask_transpose_and_accumulate(engine, T, N, ld):
    """ksa_ingest_transpose_mask: bytes [N][src_ld] -> [T][ld]; with accumulate the masks of several
    attributes are OR-ed (a failed Atoi on any attribute fails the poll); nothing outside [T][N]
    is written."""
    import torch

    rng = np.random.default_rng(T * 1000 + N)
    a = torch.from_numpy(rng.choice(np.array([0, 0, 0, 1], dtype=np.uint8), size=(N, T + 3))).to(engine.dev)
    b = torch.from_numpy(rng.choice(np.array([0, 0, 1], dtype=np.uint8), size=(N, T + 3))).to(engine.dev)
    out = torch.full((T + 2, ld), 0xEE, dtype=torch.uint8, device=engine.dev)
    engine.ingest_transpose_mask(a[:, :T], T, N, ld, out=out[1:T + 1])
    assert torch.equal(out[1:T + 1, :N], a[:, :T].t())
    engine.ingest_transpose_mask(b[:, :T], T, N, ld, out=out[1:T + 1], accumulate=True)
    assert torch.equal(out[1:T + 1, :N], (a[:, :T] | b[:, :T]).t())
    assert (out[0] == 0xEE).all() and (out[T + 1] == 0xEE).all() and (out[1:T + 1, N:] == 0xEE).all()


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", ["i64", "i32", "text16"])
@pytest.mark.parametrize("mode", ["summary", "full"])
def test_cli_replays_a_recorded_file(engine, tmp_path, fmt, mode):
    """ksa-replay --depth-file: a recording of two attributes, trace-major, as int64 / int32 / decimal
    strings (some of them unparsable: those polls fail as in sqs/sqs.go:58-61). Totals and the
    fingerprint digest equal the oracle's replay of the same values."""
    import os
    import re
    import subprocess

    from tests import oracle_lib

    rng = np.random.default_rng(len(fmt) * 10 + len(mode))
    A, N, T = 2, 333, 150
    vals = rng.integers(0, 200, size=(A, N, T)).astype(np.int64)
    fail = np.zeros((T, N), dtype=np.uint8)
    path = tmp_path / f"depth.{fmt}"
    if fmt == "i64":
        vals.tofile(path)
    elif fmt == "i32":
        vals.astype(np.int32).tofile(path)
    else:
        texts = [str(int(v)).encode() for v in vals.reshape(-1)]
        bad = rng.choice(A * N * T, size=400, replace=False)
        for i in bad:
            texts[i] = [b"", b"n/a", b"12 ", b"1_0"][int(rng.integers(0, 4))]
            a, rest = divmod(int(i), N * T)
            n, t = divmod(rest, T)
            vals[a, n, t] = 0          # what the decoder stores for a failed parse
            fail[t, n] = 1             # KSA_FAIL_POLL
        records(texts, 16).tofile(path)
    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    r = subprocess.run([cli, "--poll-period=5s", "--scale-up-cool-down=10s", "--scale-down-cool-down=20s",
                        "--scale-up-messages=250", "--scale-down-messages=40", "--max-pods=7",
                        "--attribute-names=ApproximateNumberOfMessages,ApproximateNumberOfMessagesDelayed",
                        "--traces", str(N), "--ticks", str(T), "--mode", mode, "--seed", "5",
                        "--depth-file", str(path), "--depth-format", fmt], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    m = re.search(r"scale up successful (\d+) \| scale down successful (\d+) \| at max (\d+) \| at min (\d+) \| "
                  r"up cool-down skips (\d+) \| down cool-down skips (\d+)", r.stdout)
    assert m, r.stdout
    digest = int(re.search(r"fingerprints ([0-9a-f]{16})", r.stdout).group(1), 16)
    pol = ksa.Policy(attr_mask=3, scale_up_messages=250, scale_down_messages=40, max_pods=7,
                     scale_up_cool_down=10 * 10**9, scale_down_cool_down=20 * 10**9)
    d_host = np.ascontiguousarray(vals.transpose(2, 0, 1))  # [T][A][N]
    ref = oracle_lib.replay(pol.to_c(), d_host, T, A, N, N, 0, *oracle_lib.fresh_state(5, 0, N, 1, 7),
                            want_full=False, want_summary=True, fail=fail if fmt == "text16" else None)
    want = [int(ref.summary[k].sum()) for k in ("up_applied", "down_applied", "up_noop", "down_noop", "up_cool", "down_cool")]
    assert [int(x) for x in m.groups()] == want
    assert digest == int(np.bitwise_xor.reduce(ref.summary["hash"]))
    if fmt == "text16":
        f = re.search(r"failed to get SQS messages (\d+)", r.stdout)
        assert f and int(f.group(1)) == int(ref.summary["poll_failed"].sum()) > 0


@pytest.mark.gpu
def test_cli_sweeps_a_recorded_file(engine, tmp_path):
    """ksa-replay --mode sweep --depth-file: tuning the flags on a recording. The ranking equals the
    one from the oracle's per-policy replays of the same values + oracle/policy_py.sweep_reduce."""
    import os
    import re
    import subprocess

    from oracle import policy_py
    from tests import oracle_lib

    rng = np.random.default_rng(17)
    N, T = 210, 90
    vals = np.cumsum(rng.integers(-20, 21, size=(N, T)), axis=1).clip(0, 400).astype(np.int32)
    path = tmp_path / "depth.i32"
    vals.tofile(path)
    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    r = subprocess.run([cli, "--mode", "sweep", "--attribute-names=ApproximateNumberOfMessages",
                        "--scale-up-messages=80,150", "--scale-down-messages=10,30", "--scale-up-cool-down=0s,15s",
                        "--max-pods=4,8", "--traces", str(N), "--ticks", str(T), "--seed", "2", "--cost-saturated", "9",
                        "--cost-churn", "4", "--top", "16", "--depth-file", str(path), "--depth-format", "i32"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    rows = re.findall(r"#(\d+) policy (\d+) cost (-?\d+) best-for (\d+) traces", r.stdout)
    assert len(rows) == 16, r.stdout
    S = 10**9
    pols = [ksa.Policy(attr_mask=1, scale_up_messages=um, scale_down_messages=dm, scale_up_cool_down=uc * S, max_pods=mx)
            for um in (80, 150) for dm in (10, 30) for uc in (0, 15) for mx in (4, 8)]
    d_host = np.ascontiguousarray(vals.astype(np.int64).T.reshape(T, 1, N))
    st = oracle_lib.fresh_state(2, 0, N, 1, 4)
    summ = np.stack([oracle_lib.replay(p.to_c(), d_host, T, 1, N, N, 0, *st, want_full=False, want_summary=True).summary
                     for p in pols])
    cost, best, _ = policy_py.sweep_reduce(summ, w_pod_ticks=1, w_saturated=9, w_churn=4, w_blocked=0)
    order = sorted(range(len(pols)), key=lambda p: (int(cost[p]), p))
    wins = np.bincount(best, minlength=len(pols))
    for k, (rank, pid, c, w) in enumerate(rows):
        assert (int(rank), int(pid), int(c), int(w)) == (k + 1, order[k], int(cost[order[k]]), int(wins[order[k]])), rows[k]
