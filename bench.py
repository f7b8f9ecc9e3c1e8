#!/usr/bin/env python
"""bench.py — scaling decisions/s of the policy replay on B200 (BASELINE.json's metric).

A "step" is one pass of the hot path (ksa_replay_full) over one batch of synthetic input:
BASELINE.json configs[1] — 2^20 traces x 10^4 ticks, default policy (main.go:83-91), single
attribute — per GPU (weak scaling: every rank owns its own 2^20 traces, global trace ids, no
data-path collective). Prints ONE JSON line on rank 0.

  value     device-resident throughput: all ranks' decisions / max-over-ranks CUDA-event time
  e2e       the same metric through ksa_pipeline_replay with HOST buffers (pinned), H2D + kernel +
            D2H overlapped inside the timed region
  roofline  achieved = 13 B/decision x N x T / average kernel duration (CUDA events on the
            launching stream) against the measured HBM copy peak of MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (a C restatement of the Go reference; Go cannot run here) on a
            bounded sample of the same workload, all host threads, rank 0 at N=1 only
  --impl reference  times only that CPU path (the reference arm).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "scaling decisions/sec (trace x tick)"
UNIT = "decisions/s"
N_TRACES = 1 << 20
N_TICKS = 10_000
SEED = 0x5EED0001
BYTES_PER_DECISION = 13  # 8 (int64 depth) + 4 (int32 replicas) + 1 (uint8 events); SURVEY.md §8d
WORKLOAD = "configs[1]: 2^20 traces x 10^4 ticks, default policy, 1 attribute, full trajectory out"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--traces", type=int, default=N_TRACES, help="traces per GPU (default 2^20)")
    ap.add_argument("--ticks", type=int, default=N_TICKS)
    ap.add_argument("--e2e-ticks", type=int, default=1000, help="ticks per e2e step (host buffers)")
    ap.add_argument("--e2e-chunk", type=int, default=25, help="ticks per pipeline stage")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU-baseline duration")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--variant", type=int, default=0, help="kernel variant (0 = library default)")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic():
    """dram bytes per launch from the committed ncu capture, if there is one for this workload."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(p) as f:
            return json.load(f)
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml

            self.nv = pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(0.02)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_reference_rate(policy, gen, seconds: float, steps: int = 1, warmup: int = 0):
    """Time the CPU oracle on a bounded sample of the same workload (same generator, seed, policy).
    Returns (decisions/s best-of-steps..., description). All online host threads."""
    import numpy as np

    from tests import oracle_lib

    cores = oracle_lib.online_cpus()
    n = 1 << 16
    # calibration: a short run to size the sample
    t_cal = 64
    depth = oracle_lib.gen_traces(gen.to_c(), 0, t_cal, 1, 0, n, n)
    st = oracle_lib.fresh_state(gen.seed, 0, n, 1, 5)
    t0 = time.perf_counter()
    oracle_lib.replay(policy.to_c(), depth, t_cal, 1, n, n, 0, *st)
    rate = n * t_cal / max(time.perf_counter() - t0, 1e-6)
    ticks = int(max(64, min(N_TICKS, seconds * rate / n)))
    import psutil

    avail = psutil.virtual_memory().available
    ticks = int(min(ticks, max(64, avail * 0.25 / (n * 13))))
    depth = oracle_lib.gen_traces(gen.to_c(), 0, ticks, 1, 0, n, n)
    out = (np.ones((ticks, n), dtype=np.int32), np.ones((ticks, n), dtype=np.uint8))  # pages touched
    times = []
    for it in range(max(1, warmup) + steps):  # at least one untimed pass: caches, thread start-up
        t0 = time.perf_counter()
        res = oracle_lib.replay(policy.to_c(), depth, ticks, 1, n, n, 0, *st, out=out)
        dt = time.perf_counter() - t0
        if it >= max(1, warmup):
            times.append(dt)
    sample = (f"{n} traces x {ticks} ticks of the same generator/seed/policy (traces 0..{n - 1}), "
              f"full trajectory out, {res.threads} threads")
    return n * ticks, times, cores, res.threads, sample


def run_reference(args):
    """The reference arm: the reference's own CPU implementation of the path. The Go code cannot
    be built here (no Go toolchain), so this is the C oracle (`kind: port`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import kube_sqs_autoscaler_b200 as ksa
    from kube_sqs_autoscaler_b200.engine import GenParams

    policy = ksa.Policy(attr_mask=1)
    gen = GenParams(seed=SEED)
    per_step = max(2.0, min(args.cpu_seconds, 120.0 / max(1, args.steps + args.warmup)))
    decisions, times, cores, threads, sample = cpu_reference_rate(policy, gen, per_step, args.steps, args.warmup)
    total = sum(times)
    value = decisions * len(times) / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
        "data": "synthetic", "config": {"workload": WORKLOAD, "sample_per_step": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                         "host_cores_online": cores},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import kube_sqs_autoscaler_b200 as ksa
    from kube_sqs_autoscaler_b200.engine import GenParams, HostPipeline, ReplayEngine, pinned_empty

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path "
                         "(use --impl reference for the CPU reference arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = ReplayEngine(device=local)
    if args.variant:
        eng.set_variant(args.variant, 0)

    N, T = args.traces, args.ticks
    trace0 = rank * N  # global trace ids: results do not depend on the number of GPUs
    policy = ksa.Policy(attr_mask=1)  # reference defaults, ApproximateNumberOfMessages only
    gen = GenParams(seed=SEED)        # bands around the default thresholds
    dev = eng.dev

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    depth = torch.empty((T, 1, N), dtype=torch.int64, device=dev)
    replicas = torch.empty((T, N), dtype=torch.int32, device=dev)
    events = torch.empty((T, N), dtype=torch.uint8, device=dev)
    eng.gen_traces(gen, 0, T, 1, trace0, N, N, out=depth)
    state0 = eng.new_state(N, seed=SEED, lo=1, hi=5, trace0=trace0)
    state = state0.clone()
    torch.cuda.synchronize()

    def step():
        state.replicas.copy_(state0.replicas)
        state.last_up.copy_(state0.last_up)
        state.last_down.copy_(state0.last_down)
        eng.replay_full(policy, depth, T, 1, N, N, 0, state, replicas=replicas, events=events)

    for _ in range(args.warmup):
        step()
    barrier()
    eng.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_begin.record()
    for k in range(args.steps):
        state.replicas.copy_(state0.replicas)
        state.last_up.copy_(state0.last_up)
        state.last_down.copy_(state0.last_down)
        ev[k][0].record()  # around the replay kernel only: the roofline leg
        eng.replay_full(policy, depth, T, 1, N, N, 0, state, replicas=replicas, events=events)
        ev[k][1].record()
    t_end.record()
    barrier()
    clocks = sampler.stop()
    launches = eng.launch_count()
    total_ms = t_begin.elapsed_time(t_end)
    kernel_ms = [a.elapsed_time(b) for a, b in ev]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    decisions_per_step = N * T
    value = world * decisions_per_step * args.steps / (total_ms_max / 1e3)

    # sanity inside the bench: the trajectory written by the timed launches equals the summary
    # replay's fingerprint (a kernel that skipped work would not).
    st_chk = state0.clone()
    summ = eng.replay_summary(policy, depth, T, 1, N, N, 0, st_chk)
    h = eng.traj_hash(replicas, events, T, N, N)
    # Not part of the timed region or of `value`: the other two kernels of the path on the same
    # box, for the record (summary-only replay of the same block; the configs[3] policy sweep).
    other = None
    if world == 1:
        def best_ms(fn, reps=3):
            ts = []
            for _ in range(reps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                fn()
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            return min(ts)

        sm_buf = eng.new_summary(N)

        def summary_once():
            st = state0.clone()
            eng.lib.ksa_summary_clear(sm_buf.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
            eng.replay_summary(policy, depth, T, 1, N, N, 0, st, summary=sm_buf)

        ms_sum = best_ms(summary_once)
        from tools.run_configs import sweep_grid
        from kube_sqs_autoscaler_b200.engine import policy_array
        sp, sn, stt = 4096, 4096, 4096
        parr = policy_array(sweep_grid()[:sp])
        sgen = GenParams(seed=0x5EED0004, family=ksa.GEN_WALK, high_max=500, walk_step=15)
        swalk = eng.gen_walk_init(sgen, 0, sn)
        sdepth = eng.gen_traces(sgen, 0, stt, 1, 0, sn, sn, walk=swalk)
        sr0 = eng.new_state(sn, seed=sgen.seed, lo=1, hi=5).replicas
        sout = torch.empty((sp, sn, 64), dtype=torch.uint8, device=dev)
        eng.replay_sweep(parr, sdepth, stt, 1, sn, sn, sr0, summary=sout)  # warm-up
        ms_sweep = best_ms(lambda: eng.replay_sweep(parr, sdepth, stt, 1, sn, sn, sr0, summary=sout))
        other = {
            "summary_mode": {"workload": "same block, ksa_replay_summary (8 B/decision)", "ms": ms_sum,
                             "decisions_per_s": N * T / (ms_sum * 1e-3), "GBps_algorithmic": 8.0064 * N * T / (ms_sum * 1e-3) / 1e9},
            "policy_sweep": {"workload": "configs[3]: 4096 policies x 4096 traces x 4096 ticks, summary out", "ms": ms_sweep,
                             "decisions_per_s": sp * sn * stt / (ms_sweep * 1e-3), "bound": "instruction issue"},
        }
        del sdepth, sout, sm_buf
    ok = bool(torch.equal(h.view(torch.uint8), summ.view(torch.uint8).view(N, 64)[:, 56:64].contiguous().view(-1)))
    if not ok:
        raise SystemExit("bench: full-mode trajectory fingerprint != summary-mode fingerprint")

    # ---- e2e: host buffers through the C-ABI pipeline -------------------------------------------
    e2e = None
    if not args.no_e2e:
        import psutil

        te = min(args.e2e_ticks, T)
        avail = psutil.virtual_memory().available
        while te > args.e2e_chunk and te * N * 13 * max(1, world) > 0.25 * avail:
            te //= 2
        d_host = pinned_empty((te, 1, N), np.int64)
        r_host = pinned_empty((te, N), np.int32)
        e_host = pinned_empty((te, N), np.uint8)
        torch.from_numpy(d_host).copy_(depth[:te])  # same synthetic bits, now resident on the host
        torch.cuda.synchronize()
        s_r = state0.replicas.cpu().numpy()
        s_u = state0.last_up.cpu().numpy()
        s_d = state0.last_down.cpu().numpy()
        pipe = HostPipeline(local, N, 1, min(args.e2e_chunk, te))

        n_warm = max(1, min(args.warmup, 3))
        # one fresh copy of the (in/out) carry state per call, made before the clock starts: every
        # step replays the same ticks from the same start state, as the device-resident steps do
        states = [(s_r.copy(), s_u.copy(), s_d.copy()) for _ in range(n_warm + args.steps)]

        def e2e_step(k):
            r, u, d = states[k]
            pipe.replay(policy, d_host, 0, r, u, d, r_host, e_host, None)

        for k in range(n_warm):
            e2e_step(k)
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            e2e_step(n_warm + k)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        # the e2e outputs are the same bytes the device-resident run produced
        assert np.array_equal(r_host[te - 1], replicas[te - 1].cpu().numpy()), "e2e replicas differ"
        assert np.array_equal(e_host[0], events[0].cpu().numpy()), "e2e events differ"
        e2e = {"value": world * N * te * args.steps / dt, "unit": UNIT,
               "h2d_bytes_per_step": N * te * 8 + N * 20, "d2h_bytes_per_step": N * te * 5 + N * 20,
               "ticks_per_step": te, "chunk_ticks": min(args.e2e_chunk, te),
               "api": "ksa_pipeline_replay (pinned host buffers in and out)"}
        pipe.close()
        del d_host, r_host, e_host

    # ---- CPU baseline (rank 0, N=1 only) ---------------------------------------------------------
    cpu = None
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        decisions, times, cores, threads, sample = cpu_reference_rate(policy, gen, args.cpu_seconds)
        cpu = {"value": decisions / times[0], "unit": UNIT, "cores": threads, "kind": "port",
               "sample": sample, "host_cores_online": cores,
               "note": "C restatement of the Go reference (no Go toolchain on this image)"}

    if rank == 0:
        peak, peak_src = measured_peak()
        avg_kernel_s = (sum(kernel_ms) / len(kernel_ms)) / 1e3
        achieved = BYTES_PER_DECISION * decisions_per_step / avg_kernel_s / 1e9
        traffic = recorded_traffic()
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "traces_per_gpu": N, "ticks": T, "attributes": 1,
                       "policy": "reference defaults (main.go:83-91)", "generator": "bands seed 0x5EED0001",
                       "parallelism": f"trace-sharded x{world}, no collective",
                       "l2": "inputs (84 GB) >> L2, no flush needed",
                       "kernel_variant": eng.variant_name(0, args.variant or -1)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_nominal_8TBs": achieved / 8000.0,
                         "traffic": (traffic or {}).get("dram_bytes_per_launch"),
                         "algorithmic_bytes_per_launch": BYTES_PER_DECISION * decisions_per_step,
                         "kernel_ms": sum(kernel_ms) / len(kernel_ms), "peak_source": peak_src},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": launches,
            "clocks": clocks,
            "parity_check": "traj_hash(full) == summary hash for all traces: ok",
            "other_kernels": other,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
