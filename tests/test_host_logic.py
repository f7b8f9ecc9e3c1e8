"""CPU: the host-side mirror of the reference's operator interface (flags, durations, sharding)."""
from __future__ import annotations

import argparse
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import kube_sqs_autoscaler_b200 as ksa
from kube_sqs_autoscaler_b200.policy import Policy, attr_mask_from_names, parse_duration

S = 10**9


@pytest.mark.parametrize("text,ns", [
    ("5s", 5 * S), ("30s", 30 * S), ("10s", 10 * S), ("5m", 300 * S), ("1h", 3600 * S),
    ("1.5s", 1_500_000_000), ("300ms", 300_000_000), ("2h45m", (2 * 3600 + 45 * 60) * S),
    ("-1.5h", -5400 * S), ("0", 0), ("1us", 1000), ("1µs", 1000), ("7ns", 7), (".5s", S // 2),
    ("+3s", 3 * S),
])
def test_parse_duration_go_syntax(text, ns):
    assert parse_duration(text) == ns


REJECTED_DURATIONS = ["", "5", "s", "1.s.", "5 s", " 5s", "5s ", "1d", "--1s", "+", "-", ".", ".s", "1..5s", "1e3s", "５s",
                      "9223372036854775808ns", "9223372036854775807us", "2562048h", "3000000h", "1h9223372036854775807ns",
                      "99999999999999999999s", "0.9223372036854775808e1s"]
ACCEPTED_DURATIONS = [
    ("5s", 5 * S), ("2h45m", (2 * 3600 + 45 * 60) * S), ("-1.5h", -5400 * S), ("0", 0), ("+0", 0), ("-0", 0), ("1µs", 1000),
    ("1μs", 1000), (".5s", S // 2), ("1.s", S), ("0.000000001s", 1), ("0.0000000001s", 0), ("1.0000000001s", S),
    ("9223372036854775807ns", 2**63 - 1), ("-9223372036854775808ns", -(2**63)), ("2562047h47m16.854775807s", 2**63 - 1),
    ("0.3333333333333333333333333333s", 333333333), ("1.999999999999999999999h", 7200 * S),  # Go's float64 fraction rounds up to 2h
     ("100ms50us7ns", 100_050_007),
    ("00000000000000000000005s", 5 * S),
]


@pytest.mark.parametrize("text", REJECTED_DURATIONS)
def test_parse_duration_rejects(text):
    with pytest.raises(ValueError):
        parse_duration(text)


@pytest.mark.parametrize("text,ns", ACCEPTED_DURATIONS)
def test_parse_duration_edges(text, ns):
    assert parse_duration(text) == ns


def test_cli_parses_durations_like_the_python_host(built):
    """The C++ host and the Python host restate time.ParseDuration separately (integer arithmetic with
    Go's overflow checks, the fraction through float64 as Go does): same value or same rejection on
    every case above. `ksa-replay --parse-duration=X` prints the nanoseconds and exits before any
    CUDA call."""
    import subprocess

    def cli(text):
        r = subprocess.run([built.CLI, f"--parse-duration={text}"], capture_output=True, text=True)
        return r.returncode, r.stdout.strip(), r.stderr

    for text, ns in ACCEPTED_DURATIONS + [("5m", 300 * S), ("300ms", 300_000_000), ("+3s", 3 * S)]:
        rc, out, err = cli(text)
        assert rc == 0 and int(out) == ns, (text, rc, out, err)
    for text in REJECTED_DURATIONS:
        rc, out, err = cli(text)
        assert rc == 2 and "invalid duration" in err, (text, rc, out, err)


def test_reference_command_line_parses_unchanged():
    """The README's deployment command (README.md:39-54) must parse; network flags are ignored."""
    ap = argparse.ArgumentParser()
    Policy.add_flags(ap)
    ns = ap.parse_args([
        "--sqs-queue-url=https://sqs.your_aws_region.amazonaws.com/your_aws_account_number/your_queue_name",
        "--kubernetes-deployment=your-kubernetes-deployment-name", "--kubernetes-namespace=$(K8S_NAMESPACE)",
        "--aws-region=us-west-2", "--poll-period=10s", "--scale-down-cool-down=30s",
        "--scale-up-cool-down=5m", "--scale-up-messages=20", "--scale-down-messages=10",
        "--max-pods=30", "--min-pods=1",
    ])
    p = Policy.from_args(ns)
    assert (p.poll_period, p.scale_up_cool_down, p.scale_down_cool_down) == (10 * S, 300 * S, 30 * S)
    assert (p.scale_up_messages, p.scale_down_messages, p.max_pods, p.min_pods) == (20, 10, 30, 1)
    assert p.attr_mask == 0b111 and p.num_attrs == 3


def test_defaults_are_the_reference_defaults():
    ap = argparse.ArgumentParser()
    Policy.add_flags(ap)
    assert Policy.from_args(ap.parse_args([])) == Policy()  # main.go:83-93


def test_attribute_names():
    assert attr_mask_from_names("ApproximateNumberOfMessages") == 1
    assert attr_mask_from_names(" ApproximateNumberOfMessagesNotVisible , ApproximateNumberOfMessages") == 5
    with pytest.raises(ValueError):
        attr_mask_from_names("ApproximateAgeOfOldestMessage")


def test_shard_range_partitions_exactly():
    for n_total in (0, 1, 7, 64, 2**26, 1000003):
        for world in (1, 2, 3, 4, 8):
            cuts = [ksa.shard_range(n_total, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n_total
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        ksa.shard_range(10, 2, 2)


def test_engine_refuses_to_run_without_cuda(built):
    import torch

    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU box")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ksa.ReplayEngine()


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_rank_gloo_shards_and_gathers(tmp_path):
    """world_size=2 over gloo: each rank replays its shard (with the CPU oracle standing in for the
    GPU on this box), rank 0 gathers the summaries; result equals the unsharded replay."""
    worker = os.path.join(os.path.dirname(__file__), "gloo_shard_worker.py")
    out = tmp_path / "gathered.npy"
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", env["MASTER_PORT"], worker, str(out)],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = np.load(out)
    from tests import oracle_lib
    from tests.gloo_shard_worker import GEN, N_TOTAL, POLICY, T
    depth = oracle_lib.gen_traces(GEN.to_c(), 0, T, 1, 0, N_TOTAL, N_TOTAL)
    ref = oracle_lib.replay(POLICY.to_c(), depth, T, 1, N_TOTAL, N_TOTAL, 0,
                            *oracle_lib.fresh_state(GEN.seed, 0, N_TOTAL, 1, 5),
                            want_full=False, want_summary=True)
    assert got.tobytes() == ref.summary.tobytes()


def test_two_rank_gloo_allreduces_policy_costs(tmp_path):
    """world_size=2 over gloo: per-policy cost totals of the two shards, all-reduced, equal the
    totals of the unsharded grid (the only collective of the engine: shard.allreduce_policy_costs)."""
    worker = os.path.join(os.path.dirname(__file__), "gloo_reduce_worker.py")
    out = tmp_path / "costs.npy"
    port = str(_free_port())
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=port)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", port, worker, str(out)],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    from tests.gloo_reduce_worker import N_TOTAL, shard_costs
    want, _, _ = shard_costs(0, N_TOTAL)
    assert np.array_equal(np.load(out), want)


def test_sweep_reduce_restatement_matches_a_plain_loop():
    """oracle/policy_py.sweep_reduce (numpy) against the definition written as Python loops:
    wrapping int64 cost, totals per policy, argmin per trace with ties to the lowest policy."""
    from kube_sqs_autoscaler_b200 import SUMMARY_DTYPE
    from oracle import policy_py

    rng = np.random.default_rng(11)
    P, N = 5, 9
    summ = np.zeros((P, N), dtype=SUMMARY_DTYPE)
    for f in ("up_applied", "down_applied", "up_noop", "up_cool", "down_cool"):
        summ[f] = rng.integers(0, 2**32, size=(P, N), dtype=np.uint64).astype(np.uint32)
    summ["sum_replicas"] = rng.integers(-2**62, 2**62, size=(P, N), dtype=np.int64)
    summ[3] = summ[1]  # exact ties between policies 1 and 3
    w = dict(w_pod_ticks=2**33 + 1, w_saturated=-5, w_churn=2**40, w_blocked=3)

    def wrap(x):
        x &= 2**64 - 1
        return x - 2**64 if x >= 2**63 else x

    cost = [[wrap(w["w_pod_ticks"] * int(summ["sum_replicas"][p, n]) + w["w_saturated"] * int(summ["up_noop"][p, n])
                  + w["w_churn"] * (int(summ["up_applied"][p, n]) + int(summ["down_applied"][p, n]))
                  + w["w_blocked"] * (int(summ["up_cool"][p, n]) + int(summ["down_cool"][p, n])))
             for n in range(N)] for p in range(P)]
    pc, bp, bc = policy_py.sweep_reduce(summ, **w)
    assert [int(x) for x in pc] == [wrap(sum(cost[p])) for p in range(P)]
    for n in range(N):
        col = [cost[p][n] for p in range(P)]
        assert int(bc[n]) == min(col) and int(bp[n]) == col.index(min(col))
    assert 3 not in bp.tolist()


def test_cli_keeps_the_reference_flag_surface(built):
    """ksa-replay (C++ host): reference flags parse, Go-style errors for bad values, and without
    a GPU it fails loudly instead of computing anything on the CPU."""
    import torch

    cli = os.path.join(os.path.dirname(os.path.dirname(__file__)), "kube_sqs_autoscaler_b200", "ksa-replay")
    assert os.path.exists(cli)
    r = subprocess.run([cli, "--poll-period=5x"], capture_output=True, text=True)
    assert r.returncode == 2 and 'invalid value "5x" for flag -poll-period' in r.stderr
    r = subprocess.run([cli, "--scale-up-messages", "many"], capture_output=True, text=True)
    assert r.returncode == 2 and "scale-up-messages" in r.stderr
    r = subprocess.run([cli, "--no-such-flag=1"], capture_output=True, text=True)
    assert r.returncode == 2 and "flag provided but not defined: -no-such-flag" in r.stderr
    r = subprocess.run([cli, "--attribute-names=ApproximateAgeOfOldestMessage"], capture_output=True, text=True)
    assert r.returncode == 2 and "unknown attribute name" in r.stderr
    if not torch.cuda.is_available():
        r = subprocess.run([cli, "-poll-period", "10s", "--kubernetes-deployment=x", "--sqs-queue-url=y",
                            "--aws-region=us-west-2", "--traces", "64", "--ticks", "8"],
                           capture_output=True, text=True)
        assert r.returncode == 2 and "ksa_device_count" in r.stderr  # flags parsed, then no GPU


def test_sass_counts_file_matches_the_built_library(built):
    """profiles/sass_counts.json feeds bench.py's issue-rate roofline of the policy sweep; it must be
    the count of the library that is actually built (tools/sass_report.py regenerates it)."""
    import json
    import shutil

    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    from tools import sass_report

    counts, _ = sass_report.collect()
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "sass_counts.json")) as f:
        committed = json.load(f)["instructions_per_decision"]
    assert counts == committed, "kernels changed: run python tools/sass_report.py and commit profiles/"


def test_chunk_limits_keep_stamps_inside_int32():
    """csrc/ksa_policy.cuh keeps the cool-down stamps of a call in int32: a stamp is `tick + c` with
    tick <= KSA_MAX_CHUNK_TICKS and c <= KSA_COOL_CAP (with T = 2^30 and the cap at
    2^30 + 1 the sum passed INT_MAX). The cap must still exceed every tick of a call, or a clamped
    cool-down would open its gate inside the call."""
    import re

    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(here, "kube_sqs_autoscaler_b200", "csrc", "ksa_policy.cuh")).read()
    max_ticks = re.search(r"#define KSA_MAX_CHUNK_TICKS \(\(1ll << 30\) - (\d+)\)", text)
    cap = re.search(r"#define KSA_COOL_CAP \(1 << (\d+)\)", text)
    assert max_ticks and cap, "the two limits are not spelled as this test expects"
    T, c = (1 << 30) - int(max_ticks.group(1)), 1 << int(cap.group(1))
    assert T + c <= 2**31 - 1          # the largest stamp a kernel can form
    assert c > T                       # a capped cool-down stays shut for the whole call
    assert -(T + 32) > -(2**31)        # group-relative stamps of the summary kernels (ring <= 16)
    assert 32 + c <= 2**31 - 1


def test_gpu_numa_node_lookup_reads_sysfs_and_tolerates_its_absence(tmp_path):
    """shard.host_node_of_gpu / bind_to_gpu_node: the node and CPU list come from sysfs; a platform that
    reports -1, an unknown device or a missing tree leaves the process as it is (never an error)."""
    import os

    from kube_sqs_autoscaler_b200 import shard

    assert shard._cpu_list("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert shard._cpu_list("\n") == set()
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices/system/node/node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("16-31,48-63\n")
    assert shard.host_node_of_gpu("0000:1B:00.0", str(tmp_path)) == (1, set(range(16, 32)) | set(range(48, 64)))
    (dev / "numa_node").write_text("-1\n")
    assert shard.host_node_of_gpu("0000:1b:00.0", str(tmp_path)) == (-1, set())
    assert shard.host_node_of_gpu("0000:ff:00.0", str(tmp_path)) == (-1, set())
    before = os.sched_getaffinity(0)
    info = shard.bind_to_gpu_node(0, str(tmp_path))  # no CUDA here (or no such device in the fake tree)
    assert info["bound"] is False and "why" in info
    assert os.sched_getaffinity(0) == before
