"""Second, independent restatement of the reference policy in pure Python. TEST INFRASTRUCTURE.

Only tests/ may import this (small cases; pure-Python loops). It is written from the Go text, not
from oracle/ksa_oracle.c, and keeps the reference's object shape: a `SqsClient` whose
`num_messages()` sums the selected attributes (sqs/sqs.go:45-67), a `PodAutoScaler` with
`scale_up()` / `scale_down()` (scale/scale.go:54-107) working on a fake Deployment, and `run()` —
`Run` of main.go:35-80 with a virtual wall clock in integer nanoseconds (Python ints do not wrap,
so Go's wrapping is applied explicitly with `_i32` / `_i64`).

Parity status: pinned by the reference's 11 known-answer tests (tests/golden/reference_kats.json);
anything those do not exercise is "parity unpinned by the reference" (see oracle/ksa_oracle.h).
"""
from __future__ import annotations

from dataclasses import dataclass, field

EV_UP_APPLIED, EV_DOWN_APPLIED, EV_UP_NOOP, EV_DOWN_NOOP, EV_UP_COOL, EV_DOWN_COOL = 1, 2, 4, 8, 16, 32
EV_UP_FAILED, EV_DOWN_FAILED, EV_POLL_FAILED = 0x40, 0x80, 0xC0
FAIL_POLL, FAIL_UP, FAIL_DOWN = 1, 2, 4


class InjectedError(Exception):
    """Stands for the `err != nil` returns of the AWS / Kubernetes clients (sqs/sqs.go:52-54,
    scale/scale.go:56-58,73-75,83-85,101-103)."""

DEFAULT_ATTRIBUTE_NAMES = (  # sqs/sqs.go:28-32
    "ApproximateNumberOfMessages",
    "ApproximateNumberOfMessagesDelayed",
    "ApproximateNumberOfMessagesNotVisible",
)


def _i32(x: int) -> int:
    """Go int32(x): low 32 bits, two's complement."""
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x & 0x80000000 else x


def _i64(x: int) -> int:
    """Go int (64-bit) wrap-around."""
    x &= 0xFFFFFFFFFFFFFFFF
    return x - (1 << 64) if x & (1 << 63) else x


class MockSQS:
    """main_test.go:273-286 — returns whatever attributes were last set."""

    def __init__(self, attributes: dict[str, str]):
        self.attributes = dict(attributes)
        self.fail_next = False  # test hook: make the next GetQueueAttributes return an error

    def get_queue_attributes(self) -> dict[str, str]:
        if self.fail_next:
            self.fail_next = False
            raise InjectedError("Failed to get messages in SQS")  # sqs/sqs.go:52-54
        return self.attributes

    def set_queue_attributes(self, attributes: dict[str, str]) -> None:
        self.attributes = dict(attributes)


@dataclass
class SqsClient:
    """sqs/sqs.go:20-24."""

    client: MockSQS
    attribute_names: tuple[str, ...] = DEFAULT_ATTRIBUTE_NAMES

    def num_messages(self) -> int:  # sqs/sqs.go:45-67
        out = self.client.get_queue_attributes()  # :51
        messages = 0  # :56
        for attr in self.attribute_names:  # :57
            messages = _i64(messages + int(out[attr]))  # :58 strconv.Atoi, :63
        return messages


@dataclass
class FakeDeployment:
    """The one field of the fake clientset's Deployment the path touches (main_test.go:241-271)."""

    replicas: int


@dataclass
class PodAutoScaler:
    """scale/scale.go:21-29; `deployment` stands in for Client.Get/Update on the fake clientset."""

    max: int
    min: int
    scale_up_pods: int
    scale_down_pods: int
    deployment: FakeDeployment = field(default_factory=lambda: FakeDeployment(0))
    log: list[str] = field(default_factory=list)
    text: list[str] = field(default_factory=list)  # the messages of the log lines, as logrus gets them
    fail_next_up: bool = False    # test hooks: the next Get/Update of that call returns an error
    fail_next_down: bool = False

    def scale_up(self) -> None:  # scale/scale.go:54-79; returning None == returning nil
        if self.fail_next_up:
            self.fail_next_up = False
            raise InjectedError("Failed to get deployment from kube server, no scale up occurred")  # :55-58
        current = self.deployment.replicas  # :60
        if current >= _i32(self.max):  # :62
            self.log.append("max")  # :63
            self.text.append(f"More than max pods running. No scale up. Replicas: {current}")
            return
        nxt = _i32(current + _i32(self.scale_up_pods))  # :66
        if nxt > _i32(self.max):  # :67
            nxt = _i32(self.max)  # :68
        self.deployment.replicas = nxt  # :70-72
        self.log.append("up")  # :77
        self.text.append(f"Scale up successful. Replicas: {nxt}")

    def scale_down(self) -> None:  # scale/scale.go:81-107
        if self.fail_next_down:
            self.fail_next_down = False
            raise InjectedError("Failed to get deployment from kube server, no scale down occurred")  # :82-85
        current = self.deployment.replicas  # :87
        if current <= _i32(self.min):  # :89
            self.log.append("min")  # :90
            self.text.append(f"Less than min pods running. No scale down. Replicas: {current}")
            return
        nxt = _i32(current - _i32(self.scale_down_pods))  # :94
        if nxt < _i32(self.min):  # :95
            nxt = _i32(self.min)  # :96
        self.deployment.replicas = nxt  # :98-100
        self.log.append("down")  # :105
        self.text.append(f"Scale down successful. Replicas: {nxt}")


@dataclass
class Flags:
    """The flag globals of main.go:16-25 with the defaults of main.go:83-91 (durations in ns)."""

    poll_interval: int = 5_000_000_000
    scale_down_cool_period: int = 30_000_000_000
    scale_up_cool_period: int = 10_000_000_000
    scale_up_messages: int = 100
    scale_down_messages: int = 10


_LOG_TO_EVENT = {"up": EV_UP_APPLIED, "down": EV_DOWN_APPLIED, "max": EV_UP_NOOP, "min": EV_DOWN_NOOP}


def run(flags: Flags, p: PodAutoScaler, sqs: SqsClient, depth_feed, ticks: int,
        start_tick: int = 0, last_up_tick: int = 0, last_down_tick: int = 0, fail_feed=None,
        text_log: list | None = None):
    """main.go:35-80 for `ticks` iterations of the `for` loop, virtual clock.

    `depth_feed(k)` is called before tick k's poll and may call `sqs.client.set_queue_attributes`
    (what the reference's tests do once, main_test.go:49). Returns (replicas, events,
    last_up_tick, last_down_tick): per-tick lists plus the final stamps as tick indices.
    `text_log`, if given, receives one list per tick with the messages the reference logs in that
    tick (main.go:45,49,53,58,67,72; scale/scale.go:63,77,90,105), errors without their `%v` detail.
    """
    poll = flags.poll_interval
    last_scale_up_time = last_up_tick * poll  # main.go:37 — time.Now() at start == tick 0
    last_scale_down_time = last_down_tick * poll  # main.go:38
    replicas, events = [], []
    for k in range(start_tick + 1, start_tick + ticks + 1):
        now = k * poll  # main.go:41 — time.Sleep(pollInterval)
        depth_feed(k)
        ev = 0
        p.log.clear()
        p.text = lines = []
        if text_log is not None:
            text_log.append(lines)
        f = fail_feed(k) if fail_feed else 0
        sqs.client.fail_next = bool(f & FAIL_POLL)
        p.fail_next_up = bool(f & FAIL_UP)      # consumed only if ScaleUp is actually called this tick
        p.fail_next_down = bool(f & FAIL_DOWN)

        def emit(ev):
            for line in p.log:  # the log lines of scale.go:63,77,90,105 printed this tick
                ev |= _LOG_TO_EVENT[line]
            replicas.append(p.deployment.replicas)
            events.append(ev)

        try:
            num_messages = sqs.num_messages()  # main.go:43
        except InjectedError:  # main.go:44-47
            lines.append("Failed to get SQS messages")  # :45
            replicas.append(p.deployment.replicas)
            events.append(EV_POLL_FAILED)
            continue
        lines.append(f"Found {num_messages} messages in the queue")  # main.go:49
        if num_messages >= flags.scale_up_messages:  # main.go:51
            if last_scale_up_time + flags.scale_up_cool_period > now:  # :52 Add(..).After(now)
                ev |= EV_UP_COOL  # :53
                lines.append("Waiting for cool down, skipping scale up ")  # (sic: trailing space)
                emit(ev)
                continue  # :54
            try:
                p.scale_up()  # :57
            except InjectedError:  # :57-60: log, continue; no stamp
                lines.append("Failed scaling up")  # :58
                emit(ev | EV_UP_FAILED)
                continue
            last_scale_up_time = now  # :62
            last_up_tick = k
        if num_messages <= flags.scale_down_messages:  # main.go:65
            if last_scale_down_time + flags.scale_down_cool_period > now:  # :66
                ev |= EV_DOWN_COOL  # :67
                lines.append("Waiting for cool down, skipping scale down")
                emit(ev)
                continue  # :68
            try:
                p.scale_down()  # :71
            except InjectedError:  # :71-74
                lines.append("Failed scaling down")  # :72
                emit(ev | EV_DOWN_FAILED)
                continue
            last_scale_down_time = now  # :76
            last_down_tick = k
        emit(ev)
    return replicas, events, last_up_tick, last_down_tick


def replay_trace(policy: dict, depth_rows, replicas0: int, start_tick: int = 0,
                 last_up_tick: int = 0, last_down_tick: int = 0, fail_rows=None, text_log=None):
    """Convenience wrapper: `policy` uses the C-ABI field names (include/ksa_replay.h),
    `depth_rows` is a list of per-tick tuples of the selected attributes' values."""
    names = DEFAULT_ATTRIBUTE_NAMES[: len(depth_rows[0])] if depth_rows else DEFAULT_ATTRIBUTE_NAMES
    mock = MockSQS({n: "0" for n in names})
    sqs = SqsClient(mock, tuple(names))
    p = PodAutoScaler(policy["max_pods"], policy["min_pods"], policy["scale_up_pods"],
                      policy["scale_down_pods"], FakeDeployment(replicas0))
    flags = Flags(policy["poll_period_ns"], policy["scale_down_cool_down_ns"],
                  policy["scale_up_cool_down_ns"], policy["scale_up_messages"],
                  policy["scale_down_messages"])

    def feed(k):
        row = depth_rows[k - start_tick - 1]
        mock.set_queue_attributes({n: str(v) for n, v in zip(names, row)})

    fail_feed = (lambda k: fail_rows[k - start_tick - 1]) if fail_rows is not None else None
    return run(flags, p, sqs, feed, len(depth_rows), start_tick, last_up_tick, last_down_tick, fail_feed,
               text_log)


_ATOI_RE = None


def atoi(text: bytes):
    """Go strconv.Atoi on a byte string (sqs/sqs.go:58): returns the int, or None where Go returns an
    error (ErrSyntax / ErrRange). Independent of oracle/ksa_oracle.c: a regular expression and a
    range check on Python's unbounded ints."""
    import re

    global _ATOI_RE
    if _ATOI_RE is None:
        _ATOI_RE = re.compile(rb"[+-]?[0-9]+")
    if not _ATOI_RE.fullmatch(text):
        return None
    v = int(text.decode("ascii"))
    return v if -(1 << 63) <= v <= (1 << 63) - 1 else None


def sweep_reduce(summary, w_pod_ticks=1, w_saturated=0, w_churn=0, w_blocked=0):
    """numpy restatement of ksa_sweep_reduce (an extension: the reference has no cost model).
    summary: structured array [P][N] (SUMMARY_DTYPE). Returns (policy_cost [P] int64, best_policy [N]
    int32, best_cost [N] int64), all in wrapping int64 arithmetic, ties to the lowest policy."""
    import numpy as np

    with np.errstate(over="ignore"):
        u = lambda a: a.astype(np.uint64)  # noqa: E731
        cost = (np.uint64(w_pod_ticks % 2**64) * summary["sum_replicas"].astype(np.int64).view(np.uint64)
                + np.uint64(w_saturated % 2**64) * u(summary["up_noop"])
                + np.uint64(w_churn % 2**64) * (u(summary["up_applied"]) + u(summary["down_applied"]))
                + np.uint64(w_blocked % 2**64) * (u(summary["up_cool"]) + u(summary["down_cool"])))
        policy_cost = cost.sum(axis=1, dtype=np.uint64).view(np.int64)
    signed = cost.view(np.int64)
    best_policy = signed.argmin(axis=0).astype(np.int32)  # numpy argmin returns the first minimum
    best_cost = signed.min(axis=0)
    return policy_cost, best_policy, best_cost
