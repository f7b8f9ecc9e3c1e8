"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol the header declares.
No compute calls here (no GPU)."""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

from kube_sqs_autoscaler_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ksa_replay.h")


def declared_symbols() -> list[str]:
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ksa_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_what_python_binds():
    assert declared_symbols() == sorted(_abi.SIGNATURES)


def test_library_exports_every_declared_symbol(built):
    lib = C.CDLL(_abi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), f"libksareplay.so does not export {name}"
    out = subprocess.run(["nm", "-D", "--defined-only", _abi.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (ksa_[a-z0-9_]+)\b", out))
    assert set(declared_symbols()) <= exported


def test_abi_version_and_struct_sizes(built):
    lib = _abi.load()
    assert lib.ksa_abi_version() == _abi.KSA_ABI_VERSION
    assert C.sizeof(_abi.KsaPolicy) == 80
    assert C.sizeof(_abi.KsaSummary) == 64
    assert C.sizeof(_abi.KsaGenParams) == 56
    assert C.sizeof(_abi.KsaState) == 24


def test_policy_helpers_need_no_gpu(built):
    lib = _abi.load()
    p = _abi.KsaPolicy()
    assert lib.ksa_policy_default(C.byref(p)) == 0
    # main.go:83-93
    assert (p.scale_up_messages, p.scale_down_messages, p.max_pods, p.min_pods) == (100, 10, 5, 1)
    assert (p.scale_up_pods, p.scale_down_pods, p.attr_mask) == (1, 1, 7)
    assert (p.poll_period_ns, p.scale_up_cool_down_ns, p.scale_down_cool_down_ns) == (5 * 10**9, 10**10, 3 * 10**10)
    up, down = C.c_int64(), C.c_int64()
    assert lib.ksa_policy_cool_ticks(C.byref(p), C.byref(up), C.byref(down)) == 0
    assert (up.value, down.value) == (2, 6)
    p.poll_period_ns = 0
    assert lib.ksa_policy_cool_ticks(C.byref(p), C.byref(up), C.byref(down)) == _abi.KSA_EINVAL
    assert b"poll_period_ns" in lib.ksa_last_error()


def test_sass_is_sm100a_with_wide_loads(built):
    """The hot kernels are compiled for sm_100a and use 256-bit global loads."""
    obj = os.path.join(ROOT, "kube_sqs_autoscaler_b200", "csrc", "ksa_replay.o")
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert re.search(r"LDG\.E\.\S*256", out), "expected LDG.E.256 in ksa_replay SASS"
