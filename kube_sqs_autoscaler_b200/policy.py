"""The operator API of the reference, kept as the configuration surface of the replay engine.

`Policy` carries the reference's flags under their own names (main.go:83-93): the same defaults,
the same units (Go duration syntax for the three periods), the same comma-separated
`--attribute-names`. `Policy.to_c()` produces the `ksa_policy` record of include/ksa_replay.h.
Flags that only configure the network clients the replay replaces (`--aws-region`,
`--sqs-queue-url`, `--kubernetes-deployment`, `--kubernetes-namespace`; main.go:92,95-97) are
accepted and ignored so that a reference command line parses unchanged.
"""
from __future__ import annotations

import argparse
import re
from dataclasses import dataclass

from . import _abi

DEFAULT_ATTRIBUTE_NAMES = (  # sqs/sqs.go:28-32, main.go:28
    "ApproximateNumberOfMessages",
    "ApproximateNumberOfMessagesDelayed",
    "ApproximateNumberOfMessagesNotVisible",
)

_UNIT_NS = {"ns": 1, "us": 1_000, "µs": 1_000, "μs": 1_000, "ms": 1_000_000, "s": 1_000_000_000,
            "m": 60_000_000_000, "h": 3_600_000_000_000}
_LIMIT = 1 << 63


def parse_duration(text: str) -> int:
    """Go `time.ParseDuration` → nanoseconds (what flag.DurationVar does, main.go:83-85), restated
    from Go's algorithm: `[-+]?([0-9]*(\\.[0-9]*)?[a-z]+)+`, "0" allowed without a unit, no
    surrounding white space. Integer arithmetic with Go's overflow checks (a number above 2^63, a
    product or a total above 2^63 is an error; only a negative duration may be exactly 2^63 ns);
    the fraction goes through float64 as in Go: `uint64(float64(f) * (float64(unit) / scale))`.
    The C++ host (host/ksa_replay_cli.cpp) has the same function; tests compare the two.
    Raises ValueError otherwise.
    """
    bad = ValueError(f"time: invalid duration {text!r}")
    s = text
    neg = False
    if s[:1] in ("+", "-"):
        neg = s[0] == "-"
        s = s[1:]
    if s == "0":
        return 0
    if not s:
        raise bad
    d = 0
    while s:
        if not (s[0] == "." or s[0].isascii() and s[0].isdigit()):
            raise bad
        i, v = 0, 0
        while i < len(s) and s[i].isascii() and s[i].isdigit():  # leadingInt
            if v > _LIMIT // 10:
                raise bad
            v = v * 10 + int(s[i])
            if v > _LIMIT:
                raise bad
            i += 1
        pre = i > 0
        s = s[i:]
        f, scale, post = 0, 1.0, False
        if s[:1] == ".":  # leadingFraction
            s = s[1:]
            i, overflow = 0, False
            while i < len(s) and s[i].isascii() and s[i].isdigit():
                if not overflow:
                    if f > (_LIMIT - 1) // 10:
                        overflow = True
                    else:
                        y = f * 10 + int(s[i])
                        if y > _LIMIT:
                            overflow = True
                        else:
                            f, scale = y, scale * 10
                i += 1
            post = i > 0
            s = s[i:]
        if not pre and not post:
            raise bad
        i = 0
        while i < len(s) and s[i] != "." and not (s[i].isascii() and s[i].isdigit()):
            i += 1
        if i == 0:
            raise bad
        unit = _UNIT_NS.get(s[:i])
        s = s[i:]
        if unit is None or v > _LIMIT // unit:
            raise bad
        v *= unit
        if f > 0:
            v += int(float(f) * (float(unit) / scale))
            if v > _LIMIT:
                raise bad
        d += v
        if d > _LIMIT:
            raise bad
    if neg:
        return -d
    if d > _LIMIT - 1:
        raise bad
    return d


def attr_mask_from_names(names: str) -> int:
    """--attribute-names (main.go:93,103-110): comma-separated, trimmed; order is irrelevant to
    the wrapping sum (sqs/sqs.go:57-64), membership is not."""
    mask = 0
    for raw in names.split(","):
        name = raw.strip()
        if name not in DEFAULT_ATTRIBUTE_NAMES:
            raise ValueError(f"unknown attribute name {name!r}; the replay engine models "
                             f"{', '.join(DEFAULT_ATTRIBUTE_NAMES)}")
        mask |= 1 << DEFAULT_ATTRIBUTE_NAMES.index(name)
    return mask


@dataclass
class Policy:
    """One policy record. Field names are the reference's flag names with '-' → '_'."""

    poll_period: int = 5_000_000_000            # --poll-period 5s            main.go:83
    scale_down_cool_down: int = 30_000_000_000  # --scale-down-cool-down 30s  main.go:84
    scale_up_cool_down: int = 10_000_000_000    # --scale-up-cool-down 10s    main.go:85
    scale_up_messages: int = 100                # main.go:86
    scale_down_messages: int = 10               # main.go:87
    scale_up_pods: int = 1                      # main.go:88
    scale_down_pods: int = 1                    # main.go:89
    max_pods: int = 5                           # main.go:90
    min_pods: int = 1                           # main.go:91
    attr_mask: int = 0b111                      # --attribute-names, all three  main.go:93

    @property
    def num_attrs(self) -> int:
        return bin(self.attr_mask & 0b111).count("1")

    def to_c(self) -> _abi.KsaPolicy:
        return _abi.KsaPolicy(
            scale_up_messages=self.scale_up_messages,
            scale_down_messages=self.scale_down_messages,
            max_pods=self.max_pods,
            min_pods=self.min_pods,
            scale_up_pods=self.scale_up_pods,
            scale_down_pods=self.scale_down_pods,
            poll_period_ns=self.poll_period,
            scale_up_cool_down_ns=self.scale_up_cool_down,
            scale_down_cool_down_ns=self.scale_down_cool_down,
            attr_mask=self.attr_mask,
            reserved=0,
        )

    def as_abi_dict(self) -> dict:
        c = self.to_c()
        return {name: getattr(c, name) for name, _ in _abi.KsaPolicy._fields_}

    def cool_ticks(self) -> tuple[int, int]:
        """(up, down) cool-downs in ticks, computed by the library (ksa_policy_cool_ticks)."""
        import ctypes as C

        lib = _abi.load()
        up, down = C.c_int64(), C.c_int64()
        pc = self.to_c()
        _abi.check(lib.ksa_policy_cool_ticks(C.byref(pc), C.byref(up), C.byref(down)))
        return up.value, down.value

    @staticmethod
    def add_flags(ap: argparse.ArgumentParser) -> None:
        """Register the reference's flags (main.go:83-97) on an argparse parser."""
        d = Policy()
        ap.add_argument("--poll-period", default="5s")
        ap.add_argument("--scale-down-cool-down", default="30s")
        ap.add_argument("--scale-up-cool-down", default="10s")
        ap.add_argument("--scale-up-messages", type=int, default=d.scale_up_messages)
        ap.add_argument("--scale-down-messages", type=int, default=d.scale_down_messages)
        ap.add_argument("--scale-up-pods", type=int, default=d.scale_up_pods)
        ap.add_argument("--scale-down-pods", type=int, default=d.scale_down_pods)
        ap.add_argument("--max-pods", type=int, default=d.max_pods)
        ap.add_argument("--min-pods", type=int, default=d.min_pods)
        ap.add_argument("--attribute-names", default=",".join(DEFAULT_ATTRIBUTE_NAMES))
        for ignored in ("--aws-region", "--sqs-queue-url", "--kubernetes-deployment",
                        "--kubernetes-namespace"):
            ap.add_argument(ignored, default="", help="accepted for compatibility; unused")

    @staticmethod
    def from_args(ns: argparse.Namespace) -> "Policy":
        return Policy(
            poll_period=parse_duration(ns.poll_period),
            scale_down_cool_down=parse_duration(ns.scale_down_cool_down),
            scale_up_cool_down=parse_duration(ns.scale_up_cool_down),
            scale_up_messages=ns.scale_up_messages,
            scale_down_messages=ns.scale_down_messages,
            scale_up_pods=ns.scale_up_pods,
            scale_down_pods=ns.scale_down_pods,
            max_pods=ns.max_pods,
            min_pods=ns.min_pods,
            attr_mask=attr_mask_from_names(ns.attribute_names),
        )
