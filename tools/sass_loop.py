#!/usr/bin/env python
"""Instructions per decision in the steady-state loop of one kernel of libksareplay.so.

    python tools/sass_loop.py <mangled-name substring> <decisions per loop iteration>

Finds the longest backward-branch loop of the function in `cuobjdump -sass`, prints its length,
the opcode histogram and the count divided by the decisions one iteration covers."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "kube_sqs_autoscaler_b200", "libksareplay.so")
pat, per = sys.argv[1], float(sys.argv[2])
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = out.split("Function : ")
for f in funcs[1:]:
    name = f.split("\n", 1)[0].strip()
    if pat not in name:
        continue
    ins = []
    for line in f.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    best = None
    for addr, text in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", text)
        if m and int(m.group(1), 16) < addr:
            tgt = int(m.group(1), 16)
            if best is None or addr - tgt > best[1] - best[0]:
                best = (tgt, addr)
    if not best:
        print(name, "no loop")
        continue
    body = [t for a, t in ins if best[0] <= a <= best[1]]
    # the conditional flush inside the loop is skipped by a forward branch: report both
    hist = collections.Counter()
    for t in body:
        op = t.split()[1] if t.startswith("@") else t.split()[0]
        hist[op.split(".")[0]] += 1
    print(name)
    print(f"  loop 0x{best[0]:x}..0x{best[1]:x}: {len(body)} instructions, {len(body) / per:.1f} per decision (includes the rarely-taken fold/flush block)")
    print("  " + ", ".join(f"{k} {v}" for k, v in hist.most_common()))
