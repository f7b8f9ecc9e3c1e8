"""kube_sqs_autoscaler_b200 — B200-native batched replay of the kube-sqs-autoscaler scaling policy.

Only what the hot path needs: `csrc/` (CUDA kernels + the C ABI of include/ksa_replay.h, built into
libksareplay.so next to this file), the host-side mirror of the reference's operator interface
(`Policy`, `ReplayEngine`, `HostPipeline`) and the shard helpers for one-process-per-GPU runs.
Importing the package does not load CUDA; constructing an engine without the built library raises.
"""
from . import _abi
from ._abi import (EV_DOWN_APPLIED, EV_DOWN_COOL, EV_DOWN_NOOP, EV_UP_APPLIED, EV_UP_COOL,
                   EV_UP_NOOP, EVENT_NAMES, GEN_ADVERSARIAL, GEN_BANDS, GEN_WALK, SUMMARY_DTYPE,
                   KsaError)
from .policy import DEFAULT_ATTRIBUTE_NAMES, Policy, attr_mask_from_names, parse_duration
from .logs import render_log
from .shard import shard_range

__all__ = [
    "Policy", "parse_duration", "attr_mask_from_names", "DEFAULT_ATTRIBUTE_NAMES", "KsaError",
    "SUMMARY_DTYPE", "EVENT_NAMES", "GEN_BANDS", "GEN_WALK", "GEN_ADVERSARIAL", "shard_range", "render_log",
    "EV_UP_APPLIED", "EV_DOWN_APPLIED", "EV_UP_NOOP", "EV_DOWN_NOOP", "EV_UP_COOL", "EV_DOWN_COOL",
    "ReplayEngine", "HostPipeline", "GenParams", "TraceState", "pinned_empty",
]


def __getattr__(name):  # engine.py imports torch; keep `import kube_sqs_autoscaler_b200` light
    if name in ("ReplayEngine", "HostPipeline", "GenParams", "TraceState", "pinned_empty"):
        from . import engine

        return getattr(engine, name)
    raise AttributeError(name)
