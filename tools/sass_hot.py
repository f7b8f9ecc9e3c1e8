#!/usr/bin/env python
"""Instructions per decision on the hot path of a kernel's steady-state loop.

    python tools/sass_hot.py <mangled-name substring> <decisions per loop iteration> [--lib PATH] [--loop K]

The steady-state loop is the K-th backward-branch loop of the function in address order (default
0: the replay kernels' first loop; their later, longer loop is the drain loop with one guard per
tick). The hot path is the loop body without the span behind its first inner forward branch (the
rarely-taken fold / flush block of the summary kernels). Prints the count, the count per
decision and the opcode histogram per decision."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BRA = re.compile(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)")


_SASS = {}


def hot_path(so, pat, per, which=0):
    """[(kernel name, loop start, loop end, cold block or None, instructions, {opcode: count})] for
    every function of `so` whose mangled name contains `pat`."""
    if so not in _SASS:
        _SASS[so] = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    res = []
    for f in _SASS[so].split("Function : ")[1:]:
        name = f.split("\n", 1)[0].strip()
        if pat not in name:
            continue
        ins = []
        for line in f.splitlines():
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
            if m:
                ins.append((int(m.group(1), 16), m.group(2).strip()))
        loops = sorted({(int(m.group(1), 16), a) for a, t in ins for m in [BRA.search(t)] if m and int(m.group(1), 16) < a})
        if len(loops) <= which:
            continue
        lo, hi = loops[which]
        skip = None
        for a, t in ins:
            m = BRA.search(t)
            if lo <= a < hi and m and int(m.group(1), 16) > a:
                skip = (a, int(m.group(1), 16))
                break
        hot = [t for a, t in ins if lo <= a <= hi and (skip is None or a <= skip[0] or a >= skip[1])]
        hist = collections.Counter()
        for t in hot:
            tt = t.split()
            hist[(tt[1] if tt[0].startswith("@") else tt[0]).split(".")[0]] += 1
        res.append((name, lo, hi, skip, len(hot), hist))
    return res


def describe(entry, per):
    name, lo, hi, skip, n, hist = entry
    return (f"{name}\n  loop 0x{lo:x}..0x{hi:x}" + (f", cold block 0x{skip[0]:x}..0x{skip[1]:x} left out" if skip else "") +
            f": {n} instructions = {n / per:.2f} per decision\n  " +
            ", ".join(f"{k} {v / per:.2f}" for k, v in hist.most_common()))


def main():
    args = sys.argv[1:]
    so = os.path.join(ROOT, "kube_sqs_autoscaler_b200", "libksareplay.so")
    which = 0
    for flag in ("--lib", "--loop"):
        if flag in args:
            k = args.index(flag)
            if flag == "--lib":
                so = args[k + 1]
            else:
                which = int(args[k + 1])
            del args[k:k + 2]
    pat, per = args[0], float(args[1])
    found = hot_path(so, pat, per, which)
    if not found:
        print("no kernel / no such loop for", pat)
    for e in found:
        print(describe(e, per))


if __name__ == "__main__":
    main()
