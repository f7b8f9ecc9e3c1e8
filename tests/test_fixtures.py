"""tests/golden/replay_fixtures.json — the cases a Go test replays against the unmodified reference
(go/replay/reference_test.go). Here: the committed file is exactly what tools/export_fixtures.py
produces (so it cannot drift from the oracle), it is well-formed for the Go loader, and the CUDA
kernels reproduce every case through the C ABI."""
from __future__ import annotations

import json
import os

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from tests.helpers import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ("ApproximateNumberOfMessages", "ApproximateNumberOfMessagesDelayed", "ApproximateNumberOfMessagesNotVisible")


def fixture_policy(p: dict) -> ksa.Policy:
    mask = sum(1 << NAMES.index(n) for n in p["attribute_names"])
    return ksa.Policy(poll_period=p["poll_period_ns"], scale_up_cool_down=p["scale_up_cool_down_ns"],
                      scale_down_cool_down=p["scale_down_cool_down_ns"], scale_up_messages=p["scale_up_messages"],
                      scale_down_messages=p["scale_down_messages"], scale_up_pods=p["scale_up_pods"],
                      scale_down_pods=p["scale_down_pods"], max_pods=p["max_pods"], min_pods=p["min_pods"],
                      attr_mask=mask)


def test_fixture_file_is_what_the_exporter_writes():
    from tools import export_fixtures

    with open(export_fixtures.OUT) as f:
        committed = f.read()
    assert committed == export_fixtures.render(), "tests/golden/replay_fixtures.json is stale: python tools/export_fixtures.py"


def test_fixture_file_is_well_formed():
    fx = load_golden("replay_fixtures.json")
    ids = [c["id"] for c in fx["cases"]]
    assert len(ids) == len(set(ids)) and len(ids) >= 254
    assert sum(c["source"].startswith("reference test") for c in fx["cases"]) == 10  # 6 loop + 4 clamp tests
    for c in fx["cases"]:
        T = len(c["depth"])
        assert T >= 1 and len(c["replicas"]) == T and len(c["events"]) == T
        assert all(len(row) == len(c["policy"]["attribute_names"]) for row in c["depth"])
        assert all(-(2**63) <= v < 2**63 for row in c["depth"] for v in row)  # each attribute is a Go int
        assert all(-(2**31) <= r < 2**31 for r in c["replicas"])
        assert "fail" not in c or (len(c["fail"]) == T and all(0 <= f < 8 for f in c["fail"]))
        assert c["policy"]["poll_period_ns"] > 0
        json.dumps(c)


@pytest.mark.gpu
def test_kernels_reproduce_every_fixture(engine):
    """Every case through ksa_replay_full / ksa_replay_full_faults and the summary replays (trace
    replicated over a few lanes, vector and scalar kernels): per-tick replicas and events equal the
    fixture's, the summary's counters and fingerprint equal those of the fixture's trajectory."""
    import torch

    from tests import oracle_lib

    fx = load_golden("replay_fixtures.json")
    for c in fx["cases"]:
        pol = fixture_policy(c["policy"])
        rows = np.asarray(c["depth"], dtype=np.int64)
        T, A = rows.shape
        N, ld = (40, 48) if c["id"].startswith("steady-") else (35, 48)
        depth = np.zeros((T, A, ld), dtype=np.int64)
        depth[:, :, :N] = rows[:, :, None]
        fail = None
        if "fail" in c:
            fail = np.zeros((T, ld), dtype=np.uint8)
            fail[:, :N] = np.asarray(c["fail"], dtype=np.uint8)[:, None]
        d = torch.from_numpy(depth).to(engine.dev)
        f = None if fail is None else torch.from_numpy(fail).to(engine.dev)
        want_r = np.asarray(c["replicas"], dtype=np.int32)
        want_e = np.asarray(c["events"], dtype=np.uint8)
        want_h = oracle_lib.traj_hash(want_r.reshape(T, 1), want_e.reshape(T, 1), T, 1, 1)[0]
        # the long stretches also through the time-parallel paths, chunks forced (csrc/ksa_timepar.cu: closed
        # forms for saturated / idle chunks, shifts for quiet ones), N = 40 so that the 256-bit kernels apply
        chunked = c["id"].startswith("steady-")
        try:
            for scalar in (False, True, "chunks") if chunked else (False, True):
                if scalar == "chunks":
                    os.environ["KSA_TIMEPAR_CHUNKS"], os.environ["KSA_TIMEPAR_WARM"] = "5", "8"
                    scalar = False
                engine.set_variant(-2 if scalar else 0, -2 if scalar else 0)
                st = engine.state_from_host(np.full(N, c["replicas0"], dtype=np.int32))
                r, e = engine.replay_full(pol, d, T, A, N, ld, 0, st, fail=f)
                r, e = r.cpu().numpy(), e.cpu().numpy()
                assert (r[:, :N] == want_r[:, None]).all(), (c["id"], scalar)
                assert (e[:, :N] == want_e[:, None]).all(), (c["id"], scalar)
                st = engine.state_from_host(np.full(N, c["replicas0"], dtype=np.int32))
                s = engine.replay_summary(pol, d, T, A, N, ld, 0, st, fail=f).cpu().numpy().view(ksa.SUMMARY_DTYPE).reshape(-1)
                assert (s["hash"] == want_h).all() and (s["final_replicas"] == want_r[-1]).all(), (c["id"], scalar)
                assert (s["sum_replicas"] == int(want_r.astype(np.int64).sum())).all(), (c["id"], scalar)
                assert (s["up_applied"] == int(((want_e & 1) != 0).sum())).all(), (c["id"], scalar)
                assert (s["poll_failed"] == int((want_e == 0xC0).sum())).all(), (c["id"], scalar)
        finally:
            engine.set_variant(0, 0)
            os.environ.pop("KSA_TIMEPAR_CHUNKS", None)
            os.environ.pop("KSA_TIMEPAR_WARM", None)
