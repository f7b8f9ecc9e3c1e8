"""Python host over the C ABI: torch supplies device memory and streams, libksareplay does the work.

Mirrors the reference's object model for this path (same argument meaning, same error behaviour:
errors are returned/raised, nothing panics, main.go:44-47,57-60):
  Policy              <- the flag globals + scale.PodAutoScaler fields (policy.py)
  ReplayEngine.replay_full / replay_summary
                      <- Run's loop body applied to [T][A][N] depths (main.go:40-78)
  TraceState          <- Spec.Replicas + lastScaleUpTime/lastScaleDownTime per trace
Every method calls straight into the shared library; there is no torch or CPU implementation of
the policy here, and construction fails if the library is missing.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _abi
from .policy import Policy


@dataclass
class GenParams:
    """ksa_gen_params (include/ksa_replay.h)."""

    seed: int = 0x5EED0001
    family: int = _abi.GEN_BANDS
    low_max: int = 10
    high_min: int = 100
    high_max: int = 400
    walk_step: int = 15
    overflow_log2: int = 0

    def to_c(self) -> _abi.KsaGenParams:
        return _abi.KsaGenParams(seed=self.seed, family=self.family, reserved=0, low_max=self.low_max,
                                 high_min=self.high_min, high_max=self.high_max,
                                 walk_step=self.walk_step, overflow_log2=self.overflow_log2,
                                 reserved2=0)


class TraceState:
    """Per-trace carry state on the device (struct ksa_state)."""

    def __init__(self, replicas: torch.Tensor, last_up: torch.Tensor, last_down: torch.Tensor):
        assert replicas.dtype == torch.int32 and last_up.dtype == torch.int64 and last_down.dtype == torch.int64
        assert replicas.is_cuda and replicas.is_contiguous()
        self.replicas, self.last_up, self.last_down = replicas, last_up, last_down

    def to_c(self) -> _abi.KsaState:
        return _abi.KsaState(self.replicas.data_ptr(), self.last_up.data_ptr(), self.last_down.data_ptr())

    def clone(self) -> "TraceState":
        return TraceState(self.replicas.clone(), self.last_up.clone(), self.last_down.clone())

    def __len__(self) -> int:
        return self.replicas.numel()


def policy_array(policies: list[Policy]):
    """ctypes array of ksa_policy records for replay_sweep (build once, reuse across calls)."""
    return (_abi.KsaPolicy * len(policies))(*[p.to_c() for p in policies])


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


class ReplayEngine:
    """One engine per GPU (one process per GPU in multi-GPU runs; shards are independent)."""

    def __init__(self, device: int | None = None):
        self.lib = _abi.load()  # raises ImportError when the CUDA library is not built
        if not torch.cuda.is_available():
            raise RuntimeError("ReplayEngine needs a CUDA device; there is no CPU fallback")
        self.device = torch.cuda.current_device() if device is None else int(device)
        torch.cuda.set_device(self.device)
        _abi.check(self.lib.ksa_set_device(self.device))
        self.dev = torch.device("cuda", self.device)

    # ---- inputs --------------------------------------------------------------------------------
    def gen_traces(self, gen: GenParams, tick0: int, T: int, A: int, trace0: int, N: int,
                   ld: int | None = None, out: torch.Tensor | None = None,
                   walk: torch.Tensor | None = None) -> torch.Tensor:
        """Synthetic depths [T][A][ld] int64 for ticks tick0+1..tick0+T, traces trace0..trace0+N-1."""
        ld = ld or N
        if out is None:
            out = torch.zeros((T, A, ld), dtype=torch.int64, device=self.dev)
        assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() >= T * A * ld
        g = gen.to_c()
        _abi.check(self.lib.ksa_gen_traces(C.byref(g), tick0, T, A, trace0, N, ld, out.data_ptr(),
                                           walk.data_ptr() if walk is not None else None, _stream_ptr()))
        return out

    def gen_walk_init(self, gen: GenParams, trace0: int, N: int) -> torch.Tensor:
        walk = torch.empty(N, dtype=torch.int64, device=self.dev)
        g = gen.to_c()
        _abi.check(self.lib.ksa_gen_walk_init(C.byref(g), trace0, N, walk.data_ptr(), _stream_ptr()))
        return walk

    def gen_faults(self, seed: int, poll_log2: int, up_log2: int, down_log2: int, tick0: int, T: int,
                   trace0: int, N: int, ld: int | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        """Synthetic failure mask [T][ld] uint8: each FAIL_* bit with probability 2^-log2 (0 = never)."""
        ld = ld or N
        if out is None:
            out = torch.zeros((T, ld), dtype=torch.uint8, device=self.dev)
        _abi.check(self.lib.ksa_gen_faults(seed, poll_log2, up_log2, down_log2, tick0, T, trace0, N, ld,
                                           out.data_ptr(), _stream_ptr()))
        return out

    # ---- ingest -------------------------------------------------------------------------------
    def ingest_transpose(self, src: torch.Tensor, T: int, N: int, A: int = 1, a: int = 0,
                         ld: int | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        """ksa_ingest_transpose: trace-major src [N][src_ld] (int64 or int32) -> plane `a` of the
        tick-major depth array [T][A][ld] int64."""
        assert src.is_cuda and src.dim() == 2 and src.is_contiguous() and src.dtype in (torch.int64, torch.int32)
        ld = ld or N
        if out is None:
            out = torch.zeros((T, A, ld), dtype=torch.int64, device=self.dev)
        _abi.check(self.lib.ksa_ingest_transpose(src.data_ptr(), src.element_size(), src.shape[1], T, N, A, a,
                                                 ld, out.data_ptr(), _stream_ptr()))
        return out

    def ingest_decimal(self, text: torch.Tensor, want_fail: bool = True):
        """ksa_ingest_decimal: text [count][width] uint8 (NUL-padded decimal strings) -> (values int64
        [count], fail uint8 [count] with FAIL_POLL where strconv.Atoi would return an error)."""
        assert text.is_cuda and text.dtype == torch.uint8 and text.dim() == 2 and text.is_contiguous()
        count, width = text.shape
        values = torch.empty(count, dtype=torch.int64, device=self.dev)
        fail = torch.empty(count, dtype=torch.uint8, device=self.dev) if want_fail else None
        _abi.check(self.lib.ksa_ingest_decimal(text.data_ptr(), width, count, values.data_ptr(),
                                               fail.data_ptr() if fail is not None else None, _stream_ptr()))
        return values, fail

    def ingest_transpose_mask(self, src: torch.Tensor, T: int, N: int, ld: int, out: torch.Tensor | None = None,
                              accumulate: bool = False) -> torch.Tensor:
        """ksa_ingest_transpose_mask: trace-major failure bytes [N][>=T] -> tick-major [T][ld]
        (OR-ed into `out` when accumulate)."""
        assert src.is_cuda and src.dtype == torch.uint8 and src.dim() == 2 and src.stride(1) == 1
        if out is None:
            out = torch.zeros((T, ld), dtype=torch.uint8, device=self.dev)
        _abi.check(self.lib.ksa_ingest_transpose_mask(src.data_ptr(), src.stride(0), T, N, ld, out.data_ptr(),
                                                      1 if accumulate else 0, _stream_ptr()))
        return out

    def new_state(self, N: int, seed: int = 0, lo: int = 1, hi: int = 1, trace0: int = 0) -> TraceState:
        """Fresh process (main.go:37-38): replicas = lo + hash(seed, trace) mod (hi-lo+1), stamps 0."""
        st = TraceState(torch.empty(N, dtype=torch.int32, device=self.dev),
                        torch.empty(N, dtype=torch.int64, device=self.dev),
                        torch.empty(N, dtype=torch.int64, device=self.dev))
        cs = st.to_c()
        _abi.check(self.lib.ksa_state_init(seed, trace0, N, lo, hi, C.byref(cs), _stream_ptr()))
        return st

    def state_from_host(self, replicas, last_up=None, last_down=None) -> TraceState:
        r = torch.as_tensor(np.asarray(replicas, dtype=np.int32)).to(self.dev)
        n = r.numel()
        ju = torch.zeros(n, dtype=torch.int64, device=self.dev) if last_up is None else \
            torch.as_tensor(np.asarray(last_up, dtype=np.int64)).to(self.dev)
        jd = torch.zeros(n, dtype=torch.int64, device=self.dev) if last_down is None else \
            torch.as_tensor(np.asarray(last_down, dtype=np.int64)).to(self.dev)
        return TraceState(r, ju, jd)

    # ---- the hot path --------------------------------------------------------------------------
    def replay_full(self, policy: Policy, depth: torch.Tensor, T: int, A: int, N: int, ld: int,
                    tick0: int, state: TraceState, replicas: torch.Tensor | None = None,
                    events: torch.Tensor | None = None, want_replicas: bool = True,
                    want_events: bool = True, fail: torch.Tensor | None = None):
        """ksa_replay_full. Returns (replicas [T][ld] int32, events [T][ld] uint8); advances state.
        With `fail` ([T][ld] uint8 of FAIL_* bits) it is ksa_replay_full_faults: the error paths of
        Run (main.go:44-47,57-60,71-74) are replayed too."""
        assert depth.dtype == torch.int64 and depth.is_cuda and depth.is_contiguous()
        if replicas is None and want_replicas:
            replicas = torch.empty((T, ld), dtype=torch.int32, device=self.dev)
        if events is None and want_events:
            events = torch.empty((T, ld), dtype=torch.uint8, device=self.dev)
        pc, sc = policy.to_c(), state.to_c()
        rp = replicas.data_ptr() if replicas is not None else None
        ep = events.data_ptr() if events is not None else None
        if fail is None:
            _abi.check(self.lib.ksa_replay_full(C.byref(pc), depth.data_ptr(), T, A, N, ld, tick0,
                                                C.byref(sc), rp, ep, _stream_ptr()))
        else:
            assert fail.dtype == torch.uint8 and fail.is_cuda and fail.is_contiguous()
            _abi.check(self.lib.ksa_replay_full_faults(C.byref(pc), depth.data_ptr(), fail.data_ptr(), T,
                                                       A, N, ld, tick0, C.byref(sc), rp, ep, _stream_ptr()))
        return replicas, events

    def new_summary(self, N: int) -> torch.Tensor:
        s = torch.empty((N, 64), dtype=torch.uint8, device=self.dev)
        _abi.check(self.lib.ksa_summary_clear(s.data_ptr(), N, _stream_ptr()))
        return s

    def replay_summary(self, policy: Policy, depth: torch.Tensor, T: int, A: int, N: int, ld: int,
                       tick0: int, state: TraceState, summary: torch.Tensor | None = None,
                       fail: torch.Tensor | None = None) -> torch.Tensor:
        """ksa_replay_summary (ksa_replay_summary_faults when `fail` is given). `summary` ([N][64]
        uint8 = ksa_summary records) accumulates."""
        assert depth.dtype == torch.int64 and depth.is_cuda and depth.is_contiguous()
        if summary is None:
            summary = self.new_summary(N)
        pc, sc = policy.to_c(), state.to_c()
        if fail is None:
            _abi.check(self.lib.ksa_replay_summary(C.byref(pc), depth.data_ptr(), T, A, N, ld, tick0,
                                                   C.byref(sc), summary.data_ptr(), _stream_ptr()))
        else:
            assert fail.dtype == torch.uint8 and fail.is_cuda and fail.is_contiguous()
            _abi.check(self.lib.ksa_replay_summary_faults(C.byref(pc), depth.data_ptr(), fail.data_ptr(),
                                                          T, A, N, ld, tick0, C.byref(sc),
                                                          summary.data_ptr(), _stream_ptr()))
        return summary

    def replay_sweep(self, policies: list[Policy], depth: torch.Tensor, T: int, A: int, N: int, ld: int,
                     replicas0: torch.Tensor, summary: torch.Tensor | None = None) -> torch.Tensor:
        """ksa_replay_sweep: [P][N][64] uint8 summaries, every policy from a fresh process."""
        P = len(policies)
        # a prebuilt ctypes array (policy_array) skips the per-call conversion of the grid
        arr = policies if isinstance(policies, C.Array) else policy_array(policies)
        table = torch.empty(P * _abi.POLICY_K_BYTES, dtype=torch.uint8, device=self.dev)
        if summary is None:
            summary = torch.empty((P, N, 64), dtype=torch.uint8, device=self.dev)
        assert replicas0.dtype == torch.int32 and replicas0.is_cuda
        _abi.check(self.lib.ksa_replay_sweep(arr, P, table.data_ptr(), depth.data_ptr(), T, A, N, ld,
                                             replicas0.data_ptr(), summary.data_ptr(), _stream_ptr()))
        # `table` must outlive the kernel: torch's caching allocator keeps the block on this stream.
        return summary

    def replay_sweep_static(self, policies, depth: torch.Tensor, T: int, A: int, N: int, ld: int,
                            replicas0: torch.Tensor, summary: torch.Tensor) -> torch.Tensor:
        """ksa_replay_sweep_static: the sweep for hosts that cannot allocate device memory at call
        time (another framework owns the GPU's memory, or a loop that cudaMalloc/cudaFree must not
        stall). The policy table lives inside the library (at most SWEEP_STATIC_MAX_POLICIES
        policies per call), `summary` ([P][N][64] uint8) is the caller's, and this wrapper creates
        no tensor either. Calls on one GPU share that table: keep them on one stream."""
        P = len(policies)
        arr = policies if isinstance(policies, C.Array) else policy_array(policies)
        assert replicas0.dtype == torch.int32 and replicas0.is_cuda and summary.is_cuda
        _abi.check(self.lib.ksa_replay_sweep_static(arr, P, depth.data_ptr(), T, A, N, ld,
                                                    replicas0.data_ptr(), summary.data_ptr(),
                                                    _stream_ptr()))
        return summary

    def sweep_reduce(self, summary: torch.Tensor, P: int, N: int, w_pod_ticks: int = 1, w_saturated: int = 0,
                     w_churn: int = 0, w_blocked: int = 0, policy_cost: torch.Tensor | None = None):
        """ksa_sweep_reduce (extension beyond the reference): per-policy total cost [P] (accumulated
        into `policy_cost` when given, so shards add up), per-trace best policy [N] and its cost [N]."""
        if policy_cost is None:
            policy_cost = torch.zeros(P, dtype=torch.int64, device=self.dev)
        best_policy = torch.empty(N, dtype=torch.int32, device=self.dev)
        best_cost = torch.empty(N, dtype=torch.int64, device=self.dev)
        w = _abi.KsaCost(w_pod_ticks, w_saturated, w_churn, w_blocked)
        _abi.check(self.lib.ksa_sweep_reduce(summary.data_ptr(), P, N, C.byref(w), policy_cost.data_ptr(),
                                             best_policy.data_ptr(), best_cost.data_ptr(), _stream_ptr()))
        return policy_cost, best_policy, best_cost

    def traj_hash(self, replicas: torch.Tensor, events: torch.Tensor, T: int, N: int, ld: int,
                  hash_out: torch.Tensor | None = None, chain: bool = False) -> torch.Tensor:
        if hash_out is None:
            hash_out = torch.zeros(N, dtype=torch.int64, device=self.dev)
        _abi.check(self.lib.ksa_traj_hash(replicas.data_ptr(), events.data_ptr(), T, N, ld,
                                          1 if chain else 0, hash_out.data_ptr(), _stream_ptr()))
        return hash_out

    # ---- misc ----------------------------------------------------------------------------------
    def set_variant(self, full: int = 0, summary: int = 0) -> None:
        _abi.check(self.lib.ksa_set_variant(full, summary))

    def variant_name(self, mode: int, vid: int) -> str | None:
        s = self.lib.ksa_variant_name(mode, vid)
        return s.decode() if s else None

    def launch_count(self) -> int:
        return int(self.lib.ksa_launch_count())


class HostPipeline:
    """ksa_pipeline_*: replay with HOST buffers (numpy arrays, ideally pinned), copies overlapped."""

    def __init__(self, device: int, N: int, A: int, chunk_ticks: int):
        self.lib = _abi.load()
        self.handle = C.c_void_p()
        _abi.check(self.lib.ksa_pipeline_create(C.byref(self.handle), device, N, A, chunk_ticks))
        self.N, self.A = N, A

    def close(self) -> None:
        if self.handle:
            _abi.check(self.lib.ksa_pipeline_destroy(self.handle))
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def replay(self, policy: Policy, depth: np.ndarray, tick0: int, state_replicas: np.ndarray,
               state_last_up: np.ndarray, state_last_down: np.ndarray,
               replicas: np.ndarray | None = None, events: np.ndarray | None = None,
               summary: np.ndarray | None = None, fail: np.ndarray | None = None) -> None:
        """fail: optional [T][N] uint8 of KSA_FAIL_* bits -> ksa_pipeline_replay_faults."""
        T, A, N = depth.shape
        assert A == self.A and depth.dtype == np.int64 and depth.flags.c_contiguous
        for arr, dt in ((state_replicas, np.int32), (state_last_up, np.int64), (state_last_down, np.int64)):
            assert arr.dtype == dt and arr.flags.c_contiguous and arr.size == N
        if replicas is not None:
            assert replicas.dtype == np.int32 and replicas.shape == (T, N) and replicas.flags.c_contiguous
        if events is not None:
            assert events.dtype == np.uint8 and events.shape == (T, N) and events.flags.c_contiguous
        if summary is not None:
            assert summary.dtype == _abi.SUMMARY_DTYPE and summary.size == N
        pc = policy.to_c()
        ptr = lambda a: a.ctypes.data if a is not None else None  # noqa: E731
        if fail is not None:
            assert fail.dtype == np.uint8 and fail.shape == (T, N) and fail.flags.c_contiguous
            _abi.check(self.lib.ksa_pipeline_replay_faults(self.handle, C.byref(pc), ptr(depth), ptr(fail), T, N,
                                                           tick0, ptr(state_replicas), ptr(state_last_up),
                                                           ptr(state_last_down), ptr(replicas), ptr(events),
                                                           ptr(summary)))
            return
        _abi.check(self.lib.ksa_pipeline_replay(self.handle, C.byref(pc), ptr(depth), T, N, tick0,
                                                ptr(state_replicas), ptr(state_last_up),
                                                ptr(state_last_down), ptr(replicas), ptr(events),
                                                ptr(summary)))


def pinned_empty(shape, dtype) -> np.ndarray:
    """A numpy array over page-locked host memory (torch's pinned allocator), for HostPipeline."""
    t = torch.empty(int(np.prod(shape)) * np.dtype(dtype).itemsize, dtype=torch.uint8, pin_memory=True)
    arr = t.numpy().view(dtype).reshape(shape)
    return arr
