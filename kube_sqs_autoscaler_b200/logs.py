"""The replay's outputs turned back into what the live autoscaler prints.

SURVEY.md §8 a7: the event byte of a tick says which of the reference's log lines that tick
produced. `render_log` spells them out again, message for message (`main.go:45,49,53,58,67,72`,
`scale/scale.go:63,77,90,105`), so that the log of a replayed trace can be diffed against the
log of the real autoscaler run on the same queue history. Error lines carry no `%v` detail (the
replay does not know why a call failed), and logrus's time stamp / level prefix is not produced.
"""
from __future__ import annotations

from ._abi import (EV_DOWN_APPLIED, EV_DOWN_COOL, EV_DOWN_FAILED, EV_DOWN_NOOP, EV_POLL_FAILED,
                   EV_UP_APPLIED, EV_UP_COOL, EV_UP_FAILED, EV_UP_NOOP)
from .policy import Policy


def _i32(x: int) -> int:
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x & (1 << 31) else x


def render_log(policy: Policy, replicas0: int, messages, replicas, events) -> list[list[str]]:
    """One list of messages per tick for one trace.

    messages: the attribute sum of each tick (what `NumMessages` returned); replicas / events: the
    trace's column of `ksa_replay_full`'s outputs; replicas0: the replica count before the first tick.
    """
    out = []
    r_prev = int(replicas0)
    for n, r, ev in zip(messages, replicas, events):
        n, r, ev = int(n), int(r), int(ev)
        lines = []
        if ev == EV_POLL_FAILED:
            lines.append("Failed to get SQS messages")                                   # main.go:45
            out.append(lines)
            r_prev = r
            continue
        lines.append(f"Found {n} messages in the queue")                                 # main.go:49
        r_mid = r_prev  # replica count between the two halves of the tick
        if ev & EV_UP_COOL:
            lines.append("Waiting for cool down, skipping scale up ")                    # main.go:53 (sic)
        elif ev & EV_UP_FAILED:
            lines.append("Failed scaling up")                                            # main.go:58
        elif ev & EV_UP_NOOP:
            lines.append(f"More than max pods running. No scale up. Replicas: {r_prev}")  # scale.go:63
        elif ev & EV_UP_APPLIED:
            r_mid = min(_i32(r_prev + _i32(policy.scale_up_pods)), _i32(policy.max_pods))  # scale.go:66-69
            lines.append(f"Scale up successful. Replicas: {r_mid}")                      # scale.go:77
        if ev & EV_DOWN_COOL:
            lines.append("Waiting for cool down, skipping scale down")                   # main.go:67
        elif ev & EV_DOWN_FAILED:
            lines.append("Failed scaling down")                                          # main.go:72
        elif ev & EV_DOWN_NOOP:
            lines.append(f"Less than min pods running. No scale down. Replicas: {r_mid}")  # scale.go:90
        elif ev & EV_DOWN_APPLIED:
            lines.append(f"Scale down successful. Replicas: {r}")                        # scale.go:105
        out.append(lines)
        r_prev = r
    return out
