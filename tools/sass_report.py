#!/usr/bin/env python
"""Instruction counts of the default kernels' steady-state loops, from the SASS of the built library.

    python tools/sass_report.py            # rewrites profiles/sass_counts.json and profiles/r02_sass_main_loops.txt

bench.py reads profiles/sass_counts.json for the issue-rate roofline of the policy sweep
(instructions per decision x decisions/s against 148 SMs x 4 schedulers x 32 lanes x clock).
Run it again whenever a kernel or a default launch shape changes (tests/test_host_logic.py checks
that the file still matches the library when cuobjdump is available)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import sass_hot  # noqa: E402

LIB = os.path.join(ROOT, "kube_sqs_autoscaler_b200", "libksareplay.so")

# (key, what, mangled-name pattern, decisions per trip of the steady-state loop); the replay kernel's
# last template argument (Li0E) is "an ordinary call, neither chunks of the time axis nor gated" (ksa_timepar.cu)
KERNELS = [
    ("full", "ksa_replay_full, A=1, default (8 traces per thread, ring of 2)",
     "ksa_replay_kernelILi1ELi8ELi2ELi1ELi0ELi2ELi256ELi2ELi0EE", 16),
    ("full_small_n", "ksa_replay_full, A=1, below 196608 traces (1 trace per thread, ring of 8)",
     "ksa_replay_kernelILi1ELi1ELi8ELi1ELi0ELi2ELi128ELi8ELi0EE", 8),
    ("summary_ordered", "ksa_replay_summary, A=1, scale_up_messages > scale_down_messages (1 trace per thread, ring of 12)",
     "ksa_replay_kernelILi1ELi1ELi12ELi1ELi1ELi6ELi128ELi8ELi0EE", 12),
    ("summary", "ksa_replay_summary, A=1, any policy", "ksa_replay_kernelILi1ELi1ELi12ELi1ELi1ELi2ELi128ELi8ELi0EE", 12),
    ("full_faults", "ksa_replay_full_faults, A=1 (4 traces per thread, ring of 4)",
     "ksa_replay_kernelILi1ELi4ELi4ELi1ELi0ELi8ELi128ELi5ELi0EE", 16),
    ("summary_faults_ordered", "ksa_replay_summary_faults, A=1, ordered thresholds (1 trace per thread, ring of 12)",
     "ksa_replay_kernelILi1ELi1ELi12ELi1ELi1ELi9ELi128ELi7ELi0EE", 12),
    ("summary_faults", "ksa_replay_summary_faults, A=1, any policy", "ksa_replay_kernelILi1ELi1ELi12ELi1ELi1ELi10ELi128ELi7ELi0EE", 12),
    ("sweep_shared", "ksa_replay_sweep, A=1, tiles sharing their thresholds (4 policies per thread, ring of 3)",
     "ksa_sweep_kernelILi1ELi4ELi3ELi128ELi4ELi0ELi2E", 12),
    ("sweep_narrow", "ksa_replay_sweep, A=1, ordered int32 thresholds", "ksa_sweep_kernelILi1ELi4ELi3ELi128ELi4ELi0ELi1E", 12),
    ("sweep_general", "ksa_replay_sweep, A=1, any thresholds", "ksa_sweep_kernelILi1ELi4ELi3ELi128ELi4ELi0ELi0E", 12),
]


# Kernels without a steady-state loop: the whole kernel, per element it moves.
WHOLE = [
    ("transpose_int64", "ksa_ingest_transpose, int64 source, tile of 256 ticks x 32 traces, 512 threads",
     "transpose_kernelIxLi256ELi32ELi512E", 256 * 32),
    ("transpose_int32", "ksa_ingest_transpose, int32 source, tile of 64 ticks x 128 traces, 512 threads",
     "transpose_kernelIiLi64ELi128ELi512E", 64 * 128),
]


def whole_kernel(lib, pat, per):
    """(name, instructions of the whole kernel x threads per block / elements per block, histogram)."""
    import collections
    import re
    import subprocess

    if lib not in sass_hot._SASS:
        sass_hot._SASS[lib] = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    for f in sass_hot._SASS[lib].split("Function : ")[1:]:
        name = f.split("\n", 1)[0].strip()
        if pat not in name:
            continue
        hist = collections.Counter()
        for line in f.splitlines():
            m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
            if m:
                t = m.group(2).split()
                hist[(t[1] if t[0].startswith("@") else t[0]).split(".")[0]] += 1
        return name, hist
    raise SystemExit(f"no kernel matches {pat}")


def collect(lib=LIB):
    counts, text = {}, []
    for key, what, pat, per in KERNELS:
        found = sass_hot.hot_path(lib, pat, per)
        if len(found) != 1:
            raise SystemExit(f"{key}: {len(found)} kernels match {pat}: update tools/sass_report.py (a default changed?)")
        counts[key] = round(found[0][4] / per, 2)
        text.append(f"## {key}: {what}\n{sass_hot.describe(found[0], per)}\n")
    for key, what, pat, elements in WHOLE:
        name, hist = whole_kernel(lib, pat, elements)
        threads = int(pat.rsplit("ELi", 1)[1].rstrip("E"))
        n = sum(hist.values())
        counts[key] = round(n * threads / elements, 2)
        text.append(f"## {key}: {what}\n{name}\n  whole kernel, fully unrolled: {n} instructions per thread x {threads} threads / "
                    f"{elements} elements = {n * threads / elements:.2f} per element\n  " +
                    ", ".join(f"{k} {v * threads / elements:.2f}" for k, v in hist.most_common()) + "\n")
    return counts, text


def main():
    counts, text = collect()
    out = {"source": "tools/sass_report.py: cuobjdump -sass of libksareplay.so, steady-state loop without its cold fold / flush block",
           "instructions_per_decision": counts,
           "sweep_instructions_per_decision": {k[len("sweep_"):]: v for k, v in counts.items() if k.startswith("sweep_")}}
    with open(os.path.join(ROOT, "profiles", "sass_counts.json"), "w") as f:
        json.dump(out, f, indent=1)
        f.write("\n")
    with open(os.path.join(ROOT, "profiles", "r02_sass_main_loops.txt"), "w") as f:
        f.write("# Steady-state loops of the default kernels (sm_100a SASS of kube_sqs_autoscaler_b200/libksareplay.so).\n"
                "# tools/sass_report.py; opcode counts are per decision (1 decision = 1 trace x 1 tick [x 1 policy]).\n"
                "# ALU pipe (half rate): ISETP, SEL, LOP3, VIADDMNMX, VIMNMX(3), IADD3, LEA; FMA pipe: IMAD.\n\n")
        f.write("\n".join(text))
    print(json.dumps(out["instructions_per_decision"], indent=1))


if __name__ == "__main__":
    main()
