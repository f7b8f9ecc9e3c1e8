#!/usr/bin/env python
"""profiles/r02_small_n_robust.txt from the logs of tools/gpu_r02_x.sh (gpurun_out/x_{bands,walk,quiet,steady}.log):
per entry point and shape, the plain kernels against the library's own choice. Keeps the file's header (the
lines before the first '## ')."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINE = re.compile(r"^(\w+) N=\s*(\d+) T=\s*(\d+) chunks=\s*(\S+) warm=\s*\S+ launches=(\d+)\s+([\d.]+) ms .* frac=([\d.]+) same_as_plain=(\w+)")


def main():
    out_path = os.path.join(ROOT, "profiles", "r02_small_n_robust.txt")
    header = []
    for line in open(out_path):
        if line.startswith("## "):
            break
        header.append(line)
    body = []
    for fam in ("bands", "walk", "quiet", "steady"):
        rows = {}
        for line in open(os.path.join(ROOT, "gpurun_out", f"x_{fam}.log")):
            m = LINE.match(line)
            if m:
                mode, n, t, ch, launches, ms, frac, same = m.groups()
                rows.setdefault((mode, int(n), int(t)), {})[ch] = (float(ms), float(frac), int(launches), same)
        body.append(f"## {fam}\n")
        for (mode, n, t), r in rows.items():
            if "0" not in r or "auto" not in r:
                continue
            p, a = r["0"], r["auto"]
            body.append(f"{mode:15s} {n:7d} x {t:6d}   plain {p[0]:7.3f} ms ({p[1]:.3f})   auto {a[0]:7.3f} ms ({a[1]:.3f})   "
                        f"x{p[0] / a[0]:5.2f}   launches {a[2]:2d}   identical={a[3]}\n")
        body.append("\n")
    with open(out_path, "w") as f:
        f.writelines(header + body)
    sys.stdout.writelines(body)


if __name__ == "__main__":
    main()
