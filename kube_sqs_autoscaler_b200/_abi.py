"""ctypes mirror of include/ksa_replay.h and the loader of libksareplay.so.

The library is the product; there is no CPU fallback. `load()` raises if the shared object has not
been built (`python -c "import __graft_entry__ as g; g.build()"`), and every wrapper raises
`KsaError` carrying `ksa_last_error()` on a non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# KSA_LIBRARY: load another build of the same sources (the tuning build with every variant,
# libksareplay_tuning.so); default = the product library next to this file.
LIB_PATH = os.environ.get("KSA_LIBRARY") or os.path.join(HERE, "libksareplay.so")

KSA_ABI_VERSION = 1
KSA_OK, KSA_EINVAL, KSA_ECUDA, KSA_ENOMEM, KSA_ELIMIT = 0, -1, -2, -3, -4

EV_UP_APPLIED, EV_DOWN_APPLIED, EV_UP_NOOP, EV_DOWN_NOOP, EV_UP_COOL, EV_DOWN_COOL = 1, 2, 4, 8, 16, 32
EV_UP_FAILED, EV_DOWN_FAILED, EV_POLL_FAILED = 0x40, 0x80, 0xC0  # only from the *_faults entry points
FAIL_POLL, FAIL_UP, FAIL_DOWN = 1, 2, 4
EVENT_NAMES = {
    EV_UP_APPLIED: "up_applied", EV_DOWN_APPLIED: "down_applied", EV_UP_NOOP: "up_noop",
    EV_DOWN_NOOP: "down_noop", EV_UP_COOL: "up_cool", EV_DOWN_COOL: "down_cool",
}
GEN_BANDS, GEN_WALK, GEN_ADVERSARIAL = 0, 1, 2


class KsaPolicy(C.Structure):
    """struct ksa_policy (include/ksa_replay.h) — main.go:16-25,83-93 + scale/scale.go:21-29."""

    _fields_ = [
        ("scale_up_messages", C.c_int64),
        ("scale_down_messages", C.c_int64),
        ("max_pods", C.c_int64),
        ("min_pods", C.c_int64),
        ("scale_up_pods", C.c_int64),
        ("scale_down_pods", C.c_int64),
        ("poll_period_ns", C.c_int64),
        ("scale_up_cool_down_ns", C.c_int64),
        ("scale_down_cool_down_ns", C.c_int64),
        ("attr_mask", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


class KsaState(C.Structure):
    _fields_ = [
        ("replicas_dev", C.c_void_p),
        ("last_up_tick_dev", C.c_void_p),
        ("last_down_tick_dev", C.c_void_p),
    ]


class KsaSummary(C.Structure):
    _fields_ = [
        ("final_replicas", C.c_int32),
        ("min_replicas", C.c_int32),
        ("max_replicas", C.c_int32),
        ("up_applied", C.c_uint32),
        ("down_applied", C.c_uint32),
        ("up_noop", C.c_uint32),
        ("down_noop", C.c_uint32),
        ("up_cool", C.c_uint32),
        ("down_cool", C.c_uint32),
        ("poll_failed", C.c_uint32),
        ("up_failed", C.c_uint32),
        ("down_failed", C.c_uint32),
        ("sum_replicas", C.c_int64),
        ("hash", C.c_uint64),
    ]


SUMMARY_DTYPE = np.dtype(
    [
        ("final_replicas", "<i4"), ("min_replicas", "<i4"), ("max_replicas", "<i4"),
        ("up_applied", "<u4"), ("down_applied", "<u4"), ("up_noop", "<u4"), ("down_noop", "<u4"),
        ("up_cool", "<u4"), ("down_cool", "<u4"), ("poll_failed", "<u4"), ("up_failed", "<u4"),
        ("down_failed", "<u4"), ("sum_replicas", "<i8"), ("hash", "<u8"),
    ]
)
assert SUMMARY_DTYPE.itemsize == 64 == C.sizeof(KsaSummary)
POLICY_K_BYTES = 48  # sizeof(ksa_policy_k): the kernel-side policy record (csrc/ksa_policy.cuh)
SWEEP_STATIC_MAX_POLICIES = 4096  # KSA_SWEEP_STATIC_MAX_POLICIES


class KsaCost(C.Structure):
    """struct ksa_cost: weights of ksa_sweep_reduce (an extension; the reference has no cost model)."""

    _fields_ = [("w_pod_ticks", C.c_int64), ("w_saturated", C.c_int64), ("w_churn", C.c_int64),
                ("w_blocked", C.c_int64)]


class KsaGenParams(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64),
        ("family", C.c_int32),
        ("reserved", C.c_int32),
        ("low_max", C.c_int64),
        ("high_min", C.c_int64),
        ("high_max", C.c_int64),
        ("walk_step", C.c_int64),
        ("overflow_log2", C.c_uint32),
        ("reserved2", C.c_uint32),
    ]


class KsaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libksareplay error {code}: {msg}")
        self.code = code


_P = C.POINTER
_vp, _i64, _i32 = C.c_void_p, C.c_int64, C.c_int32

# name -> (restype, argtypes). Every symbol include/ksa_replay.h declares is listed here; the
# CPU-side test tests/test_abi_symbols.py checks the list against the header and the built .so.
SIGNATURES = {
    "ksa_abi_version": (C.c_int, []),
    "ksa_last_error": (C.c_char_p, []),
    "ksa_set_device": (C.c_int, [C.c_int]),
    "ksa_device_count": (C.c_int, []),
    "ksa_dev_alloc": (C.c_int, [_P(_vp), C.c_size_t]),
    "ksa_dev_free": (C.c_int, [_vp]),
    "ksa_host_alloc": (C.c_int, [_P(_vp), C.c_size_t]),
    "ksa_host_free": (C.c_int, [_vp]),
    "ksa_host_register": (C.c_int, [_vp, C.c_size_t]),
    "ksa_host_unregister": (C.c_int, [_vp]),
    "ksa_memcpy_h2d": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "ksa_memcpy_d2h": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "ksa_stream_sync": (C.c_int, [_vp]),
    "ksa_policy_default": (C.c_int, [_P(KsaPolicy)]),
    "ksa_policy_cool_ticks": (C.c_int, [_P(KsaPolicy), _P(_i64), _P(_i64)]),
    "ksa_replay_full": (C.c_int, [_P(KsaPolicy), _vp, _i64, _i32, _i64, _i64, _i64, _P(KsaState),
                                  _vp, _vp, _vp]),
    "ksa_replay_summary": (C.c_int, [_P(KsaPolicy), _vp, _i64, _i32, _i64, _i64, _i64,
                                     _P(KsaState), _vp, _vp]),
    "ksa_summary_clear": (C.c_int, [_vp, _i64, _vp]),
    "ksa_replay_full_faults": (C.c_int, [_P(KsaPolicy), _vp, _vp, _i64, _i32, _i64, _i64, _i64,
                                         _P(KsaState), _vp, _vp, _vp]),
    "ksa_replay_summary_faults": (C.c_int, [_P(KsaPolicy), _vp, _vp, _i64, _i32, _i64, _i64, _i64,
                                            _P(KsaState), _vp, _vp]),
    "ksa_gen_faults": (C.c_int, [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _i64, _i64, _i64, _i64,
                                 _i64, _vp, _vp]),
    "ksa_replay_sweep": (C.c_int, [_P(KsaPolicy), _i32, _vp, _vp, _i64, _i32, _i64, _i64, _vp,
                                   _vp, _vp]),
    "ksa_replay_sweep_static": (C.c_int, [_P(KsaPolicy), _i32, _vp, _i64, _i32, _i64, _i64, _vp,
                                          _vp, _vp]),
    "ksa_sweep_reduce": (C.c_int, [_vp, _i32, _i64, _P(KsaCost), _vp, _vp, _vp, _vp]),
    "ksa_traj_hash": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i32, _vp, _vp]),
    "ksa_gen_traces": (C.c_int, [_P(KsaGenParams), _i64, _i64, _i32, _i64, _i64, _i64, _vp, _vp,
                                 _vp]),
    "ksa_gen_walk_init": (C.c_int, [_P(KsaGenParams), _i64, _i64, _vp, _vp]),
    "ksa_state_init": (C.c_int, [C.c_uint64, _i64, _i64, _i32, _i32, _P(KsaState), _vp]),
    "ksa_ingest_transpose": (C.c_int, [_vp, _i32, _i64, _i64, _i64, _i32, _i32, _i64, _vp, _vp]),
    "ksa_ingest_decimal": (C.c_int, [_vp, _i32, _i64, _vp, _vp, _vp]),
    "ksa_ingest_transpose_mask": (C.c_int, [_vp, _i64, _i64, _i64, _i64, _vp, _i32, _vp]),
    "ksa_pipeline_create": (C.c_int, [_P(_vp), C.c_int, _i64, _i32, _i64]),
    "ksa_pipeline_destroy": (C.c_int, [_vp]),
    "ksa_pipeline_replay": (C.c_int, [_vp, _P(KsaPolicy), _vp, _i64, _i64, _i64, _vp, _vp, _vp,
                                      _vp, _vp, _vp]),
    "ksa_pipeline_replay_faults": (C.c_int, [_vp, _P(KsaPolicy), _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp,
                                             _vp, _vp, _vp]),
    "ksa_set_variant": (C.c_int, [_i32, _i32]),
    "ksa_variant_name": (C.c_char_p, [_i32, _i32]),
    "ksa_launch_count": (_i64, []),
}

_lib = None


def load(path: str | None = None) -> C.CDLL:
    """dlopen libksareplay.so and attach the prototypes. Raises if it is not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise ImportError(
            f"{p} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback."
        )
    lib = C.CDLL(p)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == the .so does not export the symbol
        fn.restype = res
        fn.argtypes = args
    got = lib.ksa_abi_version()
    if got != KSA_ABI_VERSION:
        raise ImportError(f"{p}: ABI version {got}, this package expects {KSA_ABI_VERSION}")
    if path is None:
        _lib = lib
    return lib


def check(status: int) -> None:
    if status != KSA_OK:
        msg = load().ksa_last_error()
        raise KsaError(status, msg.decode("utf-8", "replace") if msg else "")
