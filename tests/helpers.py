"""Shared helpers for the tests: fixtures -> policies, golden loaders."""
from __future__ import annotations

import json
import os

from kube_sqs_autoscaler_b200 import Policy

HERE = os.path.dirname(os.path.abspath(__file__))
S = 1_000_000_000  # one second in Go time.Duration units


# Kernel variant ids of the tables in csrc/ (mode for ksa_variant_name: 0 full, 1 summary, 2 full with
# failures, 3 summary with failures, 4 sweep launch shapes). The product build has the defaults only,
# the tuning build (python __graft_entry__.py --tuning; KSA_LIBRARY=...tuning.so) all of them.
ALL_VARIANT_IDS = {0: range(1, 64), 1: range(1, 64), 2: range(1, 32), 3: range(1, 32), 4: range(0, 64)}


def variants_present(engine, mode: int, wanted=None) -> list[int]:
    """The ids among `wanted` (default: every id) that the loaded library was built with."""
    ids = ALL_VARIANT_IDS[mode] if wanted is None else wanted
    return [i for i in ids if i == -2 or engine.variant_name(mode, i) is not None]


def load_golden(name: str):
    with open(os.path.join(HERE, "golden", name)) as f:
        return json.load(f)


def policy_from_run_kat(k: dict) -> Policy:
    return Policy(poll_period=k["poll_s"] * S, scale_up_cool_down=k["up_cool_s"] * S,
                  scale_down_cool_down=k["down_cool_s"] * S, scale_up_messages=k["up_msgs"],
                  scale_down_messages=k["down_msgs"], scale_up_pods=k["up_pods"],
                  scale_down_pods=k["down_pods"], max_pods=k["max"], min_pods=k["min"],
                  attr_mask=0b111)


def policy_from_vector(v: dict) -> Policy:
    p = v["policy"]
    return Policy(poll_period=p.get("poll_ns", 5 * S), scale_up_cool_down=p.get("up_cool_ns", 10 * S),
                  scale_down_cool_down=p.get("down_cool_ns", 30 * S),
                  scale_up_messages=p.get("up_msgs", 100), scale_down_messages=p.get("down_msgs", 10),
                  scale_up_pods=p.get("up_pods", 1), scale_down_pods=p.get("down_pods", 1),
                  max_pods=p.get("max", 5), min_pods=p.get("min", 1), attr_mask=p.get("attr_mask", 0b001))
