#!/usr/bin/env python
"""Rebuild with -Xptxas -v and print registers / spills per replay-kernel instantiation."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
r = subprocess.run([sys.executable, os.path.join(ROOT, "__graft_entry__.py"), "--force", "-v"], capture_output=True, text=True)
name = None; rows = []
for line in (r.stdout + r.stderr).splitlines():
    if "error" in line: print(line)
    m = re.search(r"Compiling entry function '(\S+)'", line)
    if m: name = m.group(1)
    m = re.search(r"(\d+) bytes spill stores", line)
    if m: spill = m.group(1)
    m = re.search(r"Used (\d+) registers", line)
    if m and name and "replay_kernel" in name:
        k = re.search(r"ILi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)E", name).groups()
        rows.append("A=%s V=%s D=%2s H=%s MODE=%s S=%s B=%3s MINB=%2s" % k + " regs %3s spill %s" % (m.group(1), spill))
filt = sys.argv[1] if len(sys.argv) > 1 else ""
print("\n".join(r for r in sorted(set(rows)) if filt in r))
