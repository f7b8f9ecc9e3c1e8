#!/usr/bin/env python
"""Write tests/golden/replay_fixtures.json: inputs and per-tick results of the policy replay, in a
form a Go test can load and check against the UNMODIFIED reference (go/replay/reference_test.go
drives scale.PodAutoScaler on the fake clientset and kubesqs.SqsClient on a mock with a virtual
clock and compares tick by tick; no CUDA and no cgo needed for that).

    python tools/export_fixtures.py            # rewrite the file
    python tools/export_fixtures.py --check    # exit 1 if the committed file differs

Cases: the reference's 6 loop tests and 4 clamp tests (tests/golden/reference_kats.json), the
hand-derived quirk vectors (tests/golden/derived_vectors.json), 200 seeded random cases, half
of them with injected call failures, and 24 long saturated / idle / quiet stretches. Every case is computed by the C oracle's literal wall-clock
restatement of Run (main.go:35-80) and, before it is written, required to agree with the oracle's
tick model and with the independent Python restatement: three readings of the Go text. What the
file cannot contain is the output of the Go code itself (no Go toolchain here): producing that
agreement is the Go test's job.

Case fields
  policy        the 9 flags of main.go:83-91 (durations in ns) and attribute_names (main.go:93)
  replicas0     Spec.Replicas of the deployment when the process starts
  depth         one row per poll: the value of each selected attribute, in attribute_names order
  fail          optional, one entry per poll: bit 0 GetQueueAttributes fails, bit 1 the Get of
                ScaleUp fails, bit 2 the Get of ScaleDown fails (if that call is reached)
  replicas      Spec.Replicas after each tick
  events        KSA_EV_* bits of each tick (include/ksa_replay.h): which log lines the tick printed
Everything is deterministic (fixed seed, sorted keys), so regenerating gives the same bytes.
"""
from __future__ import annotations

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from kube_sqs_autoscaler_b200 import Policy  # noqa: E402
from oracle import policy_py  # noqa: E402
from tests import oracle_lib  # noqa: E402
from tests.helpers import S, load_golden, policy_from_run_kat, policy_from_vector  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "replay_fixtures.json")
NAMES = policy_py.DEFAULT_ATTRIBUTE_NAMES
SEED = 20261107
N_RANDOM = 200
N_STEADY = 24


def policy_fields(pol: Policy) -> dict:
    c = pol.to_c()
    return {
        "poll_period_ns": c.poll_period_ns, "scale_up_cool_down_ns": c.scale_up_cool_down_ns,
        "scale_down_cool_down_ns": c.scale_down_cool_down_ns, "scale_up_messages": c.scale_up_messages,
        "scale_down_messages": c.scale_down_messages, "scale_up_pods": c.scale_up_pods,
        "scale_down_pods": c.scale_down_pods, "max_pods": c.max_pods, "min_pods": c.min_pods,
        "attribute_names": [NAMES[k] for k in range(3) if c.attr_mask >> k & 1],
    }


def make_case(cid: str, source: str, pol: Policy, replicas0: int, rows, fail=None) -> dict:
    """Run the three restatements on one trace and return the fixture record."""
    rows = np.asarray(rows, dtype=np.int64)
    T, A = rows.shape
    assert A == pol.num_attrs
    f = None if fail is None else np.asarray(fail, dtype=np.uint8)
    reps, evs, _, _, _ = oracle_lib.run_wallclock(pol.to_c(), rows, replicas0, fail=f)
    res = oracle_lib.replay(pol.to_c(), rows.reshape(T, A, 1), T, A, 1, 1, 0, [replicas0], [0], [0],
                            fail=None if f is None else f.reshape(T, 1))
    assert res.replicas[:, 0].tolist() == reps.tolist() and res.events[:, 0].tolist() == evs.tolist(), cid
    pr, pe, _, _ = policy_py.replay_trace(pol.as_abi_dict(), [tuple(r) for r in rows.tolist()], replicas0,
                                          fail_rows=None if f is None else f.tolist())
    assert pr == reps.tolist() and pe == evs.tolist(), cid
    case = {"id": cid, "source": source, "policy": policy_fields(pol), "replicas0": int(replicas0),
            "depth": rows.tolist(), "replicas": reps.tolist(), "events": evs.tolist()}
    if f is not None:
        case["fail"] = f.tolist()
    return case


def random_policy(rng, mask: int) -> Policy:
    poll = int(rng.choice([1, 3, S, 5 * S, 7 * S]))

    def cool():
        k = rng.integers(0, 4)
        if k == 0:
            return int(rng.integers(-2, 7)) * poll       # multiples of the poll period, also <= 0
        if k == 1:
            return int(rng.integers(0, 40 * S))           # anything: ceil(cool/poll) matters
        if k == 2:
            return int(rng.choice([0, 2**63 - 1]))
        return int(rng.integers(1, 5)) * poll + int(rng.integers(1, max(2, poll)))

    def pods():
        if rng.integers(0, 6) == 0:
            return int(rng.choice([2**31 - 1, -(2**31), 2**32 + 2, 2**40 + 7]))  # Go int -> int32 truncation
        return int(rng.integers(-3, 13))

    def msgs():
        k = rng.integers(0, 8)
        if k == 0:
            return int(rng.choice([2**63 - 1, -(2**63), 2**31, 2**32 + 3, -(2**31) - 1]))
        return int(rng.integers(-5, 600))

    return Policy(poll_period=poll, scale_up_cool_down=cool(), scale_down_cool_down=cool(),
                  scale_up_messages=msgs(), scale_down_messages=msgs(), scale_up_pods=pods(),
                  scale_down_pods=pods(), max_pods=pods(), min_pods=pods(), attr_mask=mask)


def random_rows(rng, T: int, A: int, pol: Policy):
    c = pol.to_c()
    near = np.array([c.scale_up_messages, c.scale_up_messages - 1, c.scale_up_messages + 1, c.scale_down_messages,
                     c.scale_down_messages - 1, c.scale_down_messages + 1, 0, 1, -1, 2**63 - 1, -(2**63), 2**62, 2**31],
                    dtype=object)
    rows = []
    for _ in range(T):
        kind = rng.integers(0, 4)
        if kind == 0:      # a sum that lands on or next to a threshold
            total = int(near[rng.integers(0, 6)])
            parts = [int(rng.integers(-3, 50)) for _ in range(A - 1)]
            last = total - sum(parts)
            last = (last + 2**63) % 2**64 - 2**63
            rows.append(parts + [last])
        elif kind == 1:    # extremes: the int64 sum wraps (sqs/sqs.go:63)
            rows.append([int(near[rng.integers(6, near.size)]) for _ in range(A)])
        else:
            rows.append([int(rng.integers(0, 400)) // A for _ in range(A)])
    return rows


def build() -> dict:
    kats, derived = load_golden("reference_kats.json"), load_golden("derived_vectors.json")
    cases = []
    for k in kats["run"]:
        pol = policy_from_run_kat(k)
        T = max(k["ticks"])
        c = make_case(k["name"], "reference test " + k["cite"], pol, k["init"], [k["attributes"]] * T)
        assert all(c["replicas"][t - 1] == k["expect_final"] for t in k["ticks"]), k["name"]
        cases.append(c)
    for k in kats["scale"]:
        # direct ScaleUp / ScaleDown calls (scale/scale_test.go) as a replay: no cool-downs, depth
        # forced over / under the thresholds so that each tick makes exactly the call of the test
        pol = Policy(max_pods=k["max"], min_pods=k["min"], scale_up_pods=k["up_pods"], scale_down_pods=k["down_pods"],
                     scale_up_cool_down=0, scale_down_cool_down=0, scale_up_messages=100, scale_down_messages=10,
                     attr_mask=1)
        c = make_case(k["name"], "reference test " + k["cite"], pol, k["init"],
                      [[500] if call == "up" else [0] for call in k["calls"]])
        assert c["replicas"] == k["expect"], k["name"]
        cases.append(c)
    for v in derived["vectors"]:
        c = make_case(v["id"], "derived vector (tests/golden/derived_vectors.json), not from the reference's tests",
                      policy_from_vector(v), v["r0"], v["depth"], v.get("fail"))
        assert c["replicas"] == v["replicas"] and c["events"] == v["events"], v["id"]
        cases.append(c)
    rng = np.random.default_rng(SEED)
    for i in range(N_RANDOM):
        mask = int(rng.choice([1, 2, 4, 3, 5, 6, 7]))
        pol = random_policy(rng, mask)
        if i % 4 == 0:  # an everyday policy: the reference's defaults with small variations
            pol = Policy(attr_mask=mask, scale_up_messages=int(rng.integers(50, 300)),
                         scale_down_messages=int(rng.integers(0, 40)), max_pods=int(rng.integers(3, 9)),
                         min_pods=int(rng.integers(0, 3)), scale_up_pods=int(rng.integers(1, 4)),
                         scale_down_pods=int(rng.integers(1, 4)), scale_up_cool_down=int(rng.integers(0, 5)) * 5 * S,
                         scale_down_cool_down=int(rng.integers(0, 8)) * 5 * S)
        T = int(rng.integers(1, 41))
        A = pol.num_attrs
        rows = random_rows(rng, T, A, pol)
        r0 = int(rng.choice([int(rng.integers(-3, 14)), 2**31 - 1, -(2**31), 2147483000]))
        fail = None
        if i % 2 == 1:
            fail = rng.choice(np.array([0, 0, 0, 0, 0, 1, 2, 4, 3, 6, 7], dtype=np.uint8), size=T).tolist()
        cases.append(make_case(f"random-{i:03d}", f"seeded random case (seed {SEED})", pol, r0, rows, fail))
    # Long stretches on one side of a threshold: the saturated queue (ScaleUp every cool-down period, replicas
    # at max), the idle one (ScaleDown every period, at min), the quiet one (no call at all), and saturated
    # with short dips. The scale calls of such a stretch come exactly one period apart, ceil(cool-down / poll
    # period) ticks, with the phase of the first one: what the closed forms of csrc/ksa_timepar.cu rest on,
    # so cool-downs that are not multiples of the poll period are in. (Own generator: the cases above keep
    # their bytes.)
    rng = np.random.default_rng(SEED + 1)
    for i in range(N_STEADY):
        poll = int(rng.choice([S, 5 * S, 7 * S]))
        pol = Policy(attr_mask=1, poll_period=poll, scale_up_messages=100, scale_down_messages=10,
                     max_pods=int(rng.integers(3, 9)), min_pods=int(rng.integers(0, 3)),
                     scale_up_pods=int(rng.integers(1, 4)), scale_down_pods=int(rng.integers(1, 4)),
                     scale_up_cool_down=int(rng.integers(0, 4 * poll)), scale_down_cool_down=int(rng.integers(0, 7 * poll)))
        if i % 3 == 0:  # multiples of the poll period, the usual configuration (main.go:83-91: 5 s, 10 s, 30 s)
            pol = Policy(attr_mask=1, poll_period=poll, scale_up_messages=100, scale_down_messages=10, max_pods=pol.max_pods,
                         min_pods=pol.min_pods, scale_up_pods=pol.scale_up_pods, scale_down_pods=pol.scale_down_pods,
                         scale_up_cool_down=int(rng.integers(0, 4)) * poll, scale_down_cool_down=int(rng.integers(0, 8)) * poll)
        T = int(rng.integers(60, 161))
        lead = int(rng.integers(0, 12))  # ordinary depths first: the stretch starts from some state and phase
        kind = ("saturated", "idle", "quiet", "dips")[i % 4]
        rows = [[int(rng.integers(0, 400))] for _ in range(lead)]
        for t in range(T - lead):
            if kind == "saturated":
                rows.append([int(rng.integers(100, 900))])
            elif kind == "idle":
                rows.append([int(rng.integers(0, 11))])
            elif kind == "quiet":
                rows.append([int(rng.integers(11, 100))])
            else:
                rows.append([int(rng.integers(11, 100)) if rng.integers(0, 25) == 0 else int(rng.integers(100, 900))])
        r0 = int(rng.integers(-2, 12))
        fail = None
        if i % 8 >= 6:  # a failed call takes no stamp: the next tick calls again, and the phase moves
            fail = (rng.random(T) < 0.05).astype(np.uint8) * rng.choice(np.array([1, 2, 4], dtype=np.uint8), size=T)
            fail = fail.tolist()
        cases.append(make_case(f"steady-{i:02d}-{kind}", f"seeded {kind} stretch (seed {SEED + 1})", pol, r0, rows, fail))
    return {
        "_comment": "Generated by tools/export_fixtures.py; do not edit. Inputs and expected per-tick (replicas, event) of the "
                    "scaling policy, as computed by three restatements of main.go:35-80 / scale/scale.go:54-107 / "
                    "sqs/sqs.go:45-67 that agree. go/replay/reference_test.go checks the unmodified reference against it.",
        "event_bits": {"up_applied": 1, "down_applied": 2, "up_noop": 4, "down_noop": 8, "up_cool": 16, "down_cool": 32,
                       "up_failed": 64, "down_failed": 128, "poll_failed": 192},
        "fail_bits": {"poll": 1, "scale_up": 2, "scale_down": 4},
        "cases": cases,
    }


def render() -> str:
    return json.dumps(build(), sort_keys=True, separators=(",", ":")) + "\n"


def main():
    text = render()
    if "--check" in sys.argv:
        with open(OUT) as f:
            same = f.read() == text
        print("replay_fixtures.json is", "up to date" if same else "STALE: rerun tools/export_fixtures.py")
        sys.exit(0 if same else 1)
    with open(OUT, "w") as f:
        f.write(text)
    n = len(json.loads(text)["cases"])
    print(f"wrote {OUT}: {n} cases, {len(text)} bytes")


if __name__ == "__main__":
    main()
