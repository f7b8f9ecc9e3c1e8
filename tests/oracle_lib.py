"""ctypes front-end of the CPU oracle (oracle/libksa_oracle.so). TEST INFRASTRUCTURE ONLY.

Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
Nothing under kube_sqs_autoscaler_b200/ imports it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

from kube_sqs_autoscaler_b200._abi import KsaGenParams, KsaPolicy, SUMMARY_DTYPE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "libksa_oracle.so")

_lib = None
_vp, _i64, _i32 = C.c_void_p, C.c_int64, C.c_int32


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    src = [os.path.join(ORACLE_DIR, f) for f in ("ksa_oracle.c", "ksa_oracle.h")]
    src += [os.path.join(ROOT, "include", f) for f in ("ksa_replay.h", "ksa_tracegen.h")]
    stale = not os.path.exists(ORACLE_LIB) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(ORACLE_LIB) for s in src)
    if stale:
        subprocess.run(["make", "-C", ORACLE_DIR, "-B", "libksa_oracle.so"], check=True,
                       capture_output=True)
    lib = C.CDLL(ORACLE_LIB)
    lib.ksa_oracle_num_messages.restype = _i64
    lib.ksa_oracle_num_messages.argtypes = [_vp, _i32]
    lib.ksa_oracle_scale_up.restype = C.c_uint8
    lib.ksa_oracle_scale_up.argtypes = [C.POINTER(KsaPolicy), C.POINTER(_i32)]
    lib.ksa_oracle_scale_down.restype = C.c_uint8
    lib.ksa_oracle_scale_down.argtypes = [C.POINTER(KsaPolicy), C.POINTER(_i32)]
    lib.ksa_oracle_run_wallclock.restype = None
    lib.ksa_oracle_run_wallclock.argtypes = [C.POINTER(KsaPolicy), _vp, _vp, _i64, _i32, _i64,
                                             C.POINTER(_i32), C.POINTER(_i64), C.POINTER(_i64), _vp, _vp]
    lib.ksa_oracle_cool_ticks.restype = C.c_int
    lib.ksa_oracle_cool_ticks.argtypes = [C.POINTER(KsaPolicy), C.POINTER(_i64), C.POINTER(_i64)]
    lib.ksa_oracle_replay.restype = C.c_int
    lib.ksa_oracle_replay.argtypes = [C.POINTER(KsaPolicy), _vp, _vp, _i64, _i32, _i64, _i64, _i64, _vp,
                                      _vp, _vp, _vp, _vp, _vp, C.c_int]
    lib.ksa_oracle_gen_faults.restype = None
    lib.ksa_oracle_gen_faults.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _i64, _i64,
                                          _i64, _i64, _i64, _vp]
    lib.ksa_oracle_summary_clear.restype = None
    lib.ksa_oracle_summary_clear.argtypes = [_vp, _i64]
    lib.ksa_oracle_traj_hash.restype = None
    lib.ksa_oracle_traj_hash.argtypes = [_vp, _vp, _i64, _i64, _i64, _i32, _vp]
    lib.ksa_oracle_gen_traces.restype = C.c_int
    lib.ksa_oracle_gen_traces.argtypes = [C.POINTER(KsaGenParams), _i64, _i64, _i32, _i64, _i64, _i64,
                                          _vp, _vp, C.c_int]
    lib.ksa_oracle_gen_walk_init.restype = None
    lib.ksa_oracle_gen_walk_init.argtypes = [C.POINTER(KsaGenParams), _i64, _i64, _vp]
    lib.ksa_oracle_state_init.restype = None
    lib.ksa_oracle_state_init.argtypes = [C.c_uint64, _i64, _i64, _i32, _i32, _vp, _vp, _vp]
    lib.ksa_oracle_atoi.restype = _i64
    lib.ksa_oracle_atoi.argtypes = [_vp, _i32, _i64, _vp, _vp]
    lib.ksa_oracle_online_cpus.restype = C.c_int
    lib.ksa_oracle_online_cpus.argtypes = []
    _lib = lib
    return lib


def _p(a):
    return a.ctypes.data if a is not None else None


@dataclass
class OracleResult:
    replicas: np.ndarray | None
    events: np.ndarray | None
    state_replicas: np.ndarray
    state_last_up: np.ndarray
    state_last_down: np.ndarray
    summary: np.ndarray | None
    threads: int


def online_cpus() -> int:
    return load().ksa_oracle_online_cpus()


def cool_ticks(policy: KsaPolicy) -> tuple[int, int]:
    up, down = _i64(), _i64()
    rc = load().ksa_oracle_cool_ticks(C.byref(policy), C.byref(up), C.byref(down))
    if rc != 0:
        raise ValueError("poll_period_ns <= 0")
    return up.value, down.value


def fresh_state(seed: int, trace0: int, N: int, lo: int, hi: int):
    r = np.empty(N, dtype=np.int32)
    ju = np.empty(N, dtype=np.int64)
    jd = np.empty(N, dtype=np.int64)
    load().ksa_oracle_state_init(seed, trace0, N, lo, hi, _p(r), _p(ju), _p(jd))
    return r, ju, jd


def replay(policy: KsaPolicy, depth: np.ndarray, T: int, A: int, N: int, ld: int, tick0: int,
           state_replicas: np.ndarray, state_last_up: np.ndarray, state_last_down: np.ndarray,
           want_full: bool = True, want_summary: bool = False, summary: np.ndarray | None = None,
           threads: int = 0, out: tuple[np.ndarray, np.ndarray] | None = None,
           fail: np.ndarray | None = None) -> OracleResult:
    """Batch tick-model replay. State arrays are copied, not modified in place. `out` = preallocated
    (replicas [T][ld] int32, events [T][ld] uint8) to keep page faults out of a timed call."""
    depth = np.ascontiguousarray(depth, dtype=np.int64)
    assert depth.size >= T * A * ld
    r = np.array(state_replicas, dtype=np.int32, copy=True)
    ju = np.array(state_last_up, dtype=np.int64, copy=True)
    jd = np.array(state_last_down, dtype=np.int64, copy=True)
    if out is not None:
        replicas, events = out
        assert replicas.dtype == np.int32 and events.dtype == np.uint8
        assert replicas.size >= T * ld and events.size >= T * ld
    else:
        replicas = np.zeros((T, ld), dtype=np.int32) if want_full else None
        events = np.zeros((T, ld), dtype=np.uint8) if want_full else None
    if want_summary and summary is None:
        summary = np.empty(N, dtype=SUMMARY_DTYPE)
        load().ksa_oracle_summary_clear(_p(summary), N)
    elif summary is not None:
        summary = np.array(summary, copy=True)
    if fail is not None:
        fail = np.ascontiguousarray(fail, dtype=np.uint8)
        assert fail.size >= T * ld
    used = load().ksa_oracle_replay(C.byref(policy), _p(depth), _p(fail), T, A, N, ld, tick0, _p(r), _p(ju),
                                    _p(jd), _p(replicas), _p(events), _p(summary), threads)
    if used < 0:
        raise ValueError("ksa_oracle_replay rejected its arguments")
    return OracleResult(replicas, events, r, ju, jd, summary, used)


def run_wallclock(policy: KsaPolicy, depth_rows: np.ndarray, replicas0: int, tick0: int = 0,
                  last_up: int = 0, last_down: int = 0, fail=None):
    """Literal main.go:35-80 restatement, one trace. depth_rows: [T][A]."""
    d = np.ascontiguousarray(depth_rows, dtype=np.int64)
    T, A = d.shape
    r = _i32(replicas0)
    ju, jd = _i64(last_up), _i64(last_down)
    reps = np.zeros(T, dtype=np.int32)
    evs = np.zeros(T, dtype=np.uint8)
    fa = np.ascontiguousarray(fail, dtype=np.uint8) if fail is not None else None
    load().ksa_oracle_run_wallclock(C.byref(policy), _p(d), _p(fa), T, A, tick0, C.byref(r), C.byref(ju),
                                    C.byref(jd), _p(reps), _p(evs))
    return reps, evs, r.value, ju.value, jd.value


def gen_traces(gen: KsaGenParams, tick0: int, T: int, A: int, trace0: int, N: int, ld: int,
               walk: np.ndarray | None = None, threads: int = 0) -> np.ndarray:
    depth = np.zeros((T, A, ld), dtype=np.int64)
    rc = load().ksa_oracle_gen_traces(C.byref(gen), tick0, T, A, trace0, N, ld, _p(depth), _p(walk), threads)
    if rc < 0:
        raise ValueError("ksa_oracle_gen_traces rejected its arguments")
    return depth


def gen_faults(seed: int, poll_log2: int, up_log2: int, down_log2: int, tick0: int, T: int,
               trace0: int, N: int, ld: int) -> np.ndarray:
    fail = np.zeros((T, ld), dtype=np.uint8)
    load().ksa_oracle_gen_faults(seed, poll_log2, up_log2, down_log2, tick0, T, trace0, N, ld, _p(fail))
    return fail


def gen_walk_init(gen: KsaGenParams, trace0: int, N: int) -> np.ndarray:
    walk = np.empty(N, dtype=np.int64)
    load().ksa_oracle_gen_walk_init(C.byref(gen), trace0, N, _p(walk))
    return walk


def traj_hash(replicas: np.ndarray, events: np.ndarray, T: int, N: int, ld: int,
              start: np.ndarray | None = None) -> np.ndarray:
    h = np.zeros(N, dtype=np.uint64) if start is None else np.array(start, dtype=np.uint64, copy=True)
    load().ksa_oracle_traj_hash(_p(np.ascontiguousarray(replicas)), _p(np.ascontiguousarray(events)),
                                T, N, ld, 0 if start is None else 1, _p(h))
    return h


def scale_calls(policy: KsaPolicy, init: int, calls: list[str]) -> list[int]:
    """Direct ScaleUp/ScaleDown calls, as scale/scale_test.go does."""
    r = _i32(init)
    out = []
    for c in calls:
        fn = load().ksa_oracle_scale_up if c == "up" else load().ksa_oracle_scale_down
        fn(C.byref(policy), C.byref(r))
        out.append(r.value)
    return out


def num_messages(values) -> int:
    a = np.asarray(values, dtype=np.int64)
    return load().ksa_oracle_num_messages(_p(a), a.size)


def atoi_records(text: np.ndarray):
    """text: uint8 [count][width] fixed-width NUL-padded records -> (values int64, fail uint8)."""
    text = np.ascontiguousarray(text, dtype=np.uint8)
    count, width = text.shape
    values = np.zeros(count, dtype=np.int64)
    fail = np.zeros(count, dtype=np.uint8)
    load().ksa_oracle_atoi(_p(text), width, count, _p(values), _p(fail))
    return values, fail
