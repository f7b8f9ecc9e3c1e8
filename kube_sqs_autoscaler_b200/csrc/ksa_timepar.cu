// ksa_timepar.cu — full-trajectory replay for trace counts that cannot fill the GPU: the time axis
// is cut into K chunks per trace and the chunks are replayed in parallel.
//
// One thread per trace is bound by its own dependent chain (about 41 instructions per tick, one
// after the other) once there are fewer traces than the GPU has thread slots: 4096 traces are 128
// warps on 148 SMs (profiles/r01_small_n_variants.txt: 0.10 of the copy peak). The policy's state is
// tiny — the replica count and the two cool-down stamps (main.go:37-38, scale/scale.go:60) — and a
// run started from a WRONG state soon behaves like the true one: the replica count is clamped into
// [min, max] (scale/scale.go:67-69, 95-97) and each stamp is overwritten by the next scale call
// (main.go:62, 76). Once the whole state is equal, the two runs are identical from there on (the
// machine is deterministic in its state and the inputs are the same). That gives an exact scheme:
//
//  pass 1  thread (trace, chunk): replay the chunk from a guessed entry state (chunk 0: the true one),
//          write its outputs, leave the state after the chunk and after a few checkpoint ticks in a
//          scratch record;
//  pass 2  thread (trace, chunk >= 1): take the previous chunk's pass-1 exit state as entry state. If it
//          behaves like the guess, nothing to do. Otherwise replay from it, rewriting the outputs,
//          until the state at a checkpoint equals what pass 1 left there: from that tick on the
//          pass-1 outputs are already right. If no checkpoint matches, the whole chunk is rewritten;
//  pass 3  thread (trace): walk the chunks in order with the TRUE state. A chunk whose pass-2 entry
//          state was the true one is accepted as it is (by induction all are, as long as every chunk
//          before it reached a matching checkpoint); otherwise the chunk is replayed again from the
//          true state until a checkpoint matches. Writes the carry state back.
// Results are bit-identical to the sequential replay whatever K is and whether or not any chunk
// converges (a trace that never does is simply replayed sequentially by pass 3); the tests force
// small chunks on adversarial policies (one-sided clamps with r0 outside [min, max], min > max,
// cool-downs longer than the call) to check exactly that.
//
// "Equal state" at tick t means equal replica counts and equal gates for every later tick: a stamp
// that opens its gate at or before tick t + 1 is as good as any other such stamp. The stamps
// written back to the caller are the true ones: pass 3 takes, per direction, the stamp of the last
// scale call of the true trajectory (a pass-1 stamp only counts if it was taken after the tick
// from which pass 1's trajectory is the true one).
//
// Scratch: one 128-byte record per (trace, chunk) in a fixed device array of the library image
// (kMaxChunks records, 32 MB per GPU), so the entry point still allocates nothing. One time-parallel
// replay per device at a time: a call that finds the scratch busy takes the plain kernels instead.
#include "ksa_internal.h"

#include <cstdlib>
#include <mutex>

namespace {

constexpr int kMaxChunks = 1 << 18;   // records: traces x chunks of one call
constexpr int kChecks = 8;
__host__ __device__ constexpr int check_tick(int j) { return 32 << j; }  // 32, 64, ..., 4096 ticks into the chunk
constexpr int D = 8;                  // ring depth of pass 1 (the checkpoint ticks are multiples of it)
constexpr int BLOCK = 128;

struct TpState {
  int r, su, sd;  // replicas; first local tick at which the up / down gate is open (ksa_policy.cuh)
};
struct TpChunk {
  int status;        // pass 2: 0 entry behaved like the guess; j+1 converged at checkpoint j; -1 never
  TpState exit1;     // pass 1: state after the chunk's last tick
  TpState stop2;     // pass 2: state where it stopped (the checkpoint, or the end of the chunk)
  TpState ck[kChecks];  // pass 1: state after chunk tick check_tick(j), if the chunk has it
  int pad;
};
static_assert(sizeof(TpChunk) == 128, "scratch record");

static __device__ __align__(256) TpChunk g_tp_scratch[kMaxChunks];

struct TpArgs {
  ksa_policy_k pk;
  const long long* depth;
  long long ld, tick0, N;
  int T, A, K, L;  // ticks of the call, attribute planes, chunks, ticks per chunk (a multiple of D)
  int one;
  int* st_r;
  long long* st_ju;
  long long* st_jd;
  int* replicas;
  unsigned char* events;
};

// A chunk has checkpoint j if pass 1 passes it in its steady-state loop (see tp_spec_kernel).
__device__ __forceinline__ bool has_check(int len, int j) { return check_tick(j) + D <= len; }

// Same behaviour from local tick t + 1 on.
__device__ __forceinline__ bool same_future(const TpState& a, const TpState& b, int t) {
  return a.r == b.r && max(a.su, t + 1) == max(b.su, t + 1) && max(a.sd, t + 1) == max(b.sd, t + 1);
}

__device__ __forceinline__ long long load_sum(const long long* p, long long ld, int A) {
  unsigned long long acc = (unsigned long long)__ldg(p);  // sqs/sqs.go:56-64, wrapping
  for (int a = 1; a < A; ++a) acc += (unsigned long long)__ldg(p + a * ld);
  return (long long)acc;
}

// ---- pass 1 ------------------------------------------------------------------------------------------
// (the ring holds 8 * A depth values: 64 registers per thread are enough for one attribute only)
template <int A>
__global__ void __launch_bounds__(BLOCK, A == 1 ? 8 : (A == 2 ? 5 : 4)) tp_spec_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const int k = blockIdx.y;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const ksa_step_consts c(a.pk);
  const int one = a.one;
  ksa_trace_regs s;
  s.r = a.st_r[n];
  if (k == 0) {
    s.su = ksa_local_stamp(a.pk.cu, c.cu, a.tick0, a.st_ju[n]);
    s.sd = ksa_local_stamp(a.pk.cd, c.cd, a.tick0, a.st_jd[n]);
  } else {
    s.su = s.sd = 0;  // the guess: the trace's starting replica count, both gates open
  }
  TpChunk* const rec = &g_tp_scratch[(long long)k * a.N + n];
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + (long long)t0 * row + n;
  int* rp = a.replicas + (long long)t0 * a.ld + n;
  unsigned char* ep = a.events + (long long)t0 * a.ld + n;

  long long buf[D][A];
#pragma unroll
  for (int j = 0; j < D; ++j)
    if (j < len) {
#pragma unroll
      for (int p = 0; p < A; ++p)
        asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(buf[j][p]) : "l"(dp + (long long)j * row + p * a.ld));
    }
  auto tick = [&](const int tt, const int j, const bool prefetch) {
    unsigned long long acc = (unsigned long long)buf[j][0];
#pragma unroll
    for (int p = 1; p < A; ++p) acc += (unsigned long long)buf[j][p];
    if (prefetch) {
#pragma unroll
      for (int p = 0; p < A; ++p)
        asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];"
                     : "=l"(buf[j][p]) : "l"(dp + (long long)(tt + D) * row + p * a.ld));
    }
    const unsigned ev = ksa_step_fma_byte(c, (long long)acc, t0 + tt + 1, one, s);
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(rp + (long long)tt * a.ld), "r"(s.r) : "memory");
    asm volatile("st.global.cs.b8 [%0], %1;" ::"l"(ep + (long long)tt * a.ld), "r"(ev) : "memory");
  };
  int t = 0;
#pragma unroll 1
  for (; t + 2 * D <= len; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) tick(t + j, j, true);
    const int done = t + D;  // ticks of the chunk replayed so far
    if ((done & (done - 1)) == 0 && done >= check_tick(0) && done <= check_tick(kChecks - 1))  // 32, 64, ..., 4096
      rec->ck[31 - __clz(done) - 5] = TpState{s.r, s.su, s.sd};
  }
#pragma unroll 1
  for (; t < len; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      if (t + j < len) tick(t + j, j, t + j + D < len);
  }
  rec->exit1 = TpState{s.r, s.su, s.sd};
  rec->status = 0;
}

// Replay chunk (t0, len) from state s, rewriting the outputs, in groups of D ticks with the depths
// of the next group already in flight (a plain loop waits a memory latency per tick). Stops after a
// checkpoint tick j >= first_check whose pass-1 state has the same future; returns that j, or -1
// when the chunk ends first.
template <int A>
__device__ int tp_replay_until_match(const TpArgs& a, const ksa_step_consts& c, long long n, int t0, int len,
                                     const TpChunk* rec, int first_check, ksa_trace_regs& s) {
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + (long long)t0 * row + n;
  int* rp = a.replicas + (long long)t0 * a.ld + n;
  unsigned char* ep = a.events + (long long)t0 * a.ld + n;
  int next = (first_check < kChecks && has_check(len, first_check)) ? first_check : kChecks;
  long long buf[D];
#pragma unroll
  for (int j = 0; j < D; ++j) buf[j] = j < len ? load_sum(dp + (long long)j * row, a.ld, A) : 0;
  for (int t = 0; t < len; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const int tt = t + j;
      if (tt < len) {
        const long long msgs = buf[j];
        if (tt + D < len) buf[j] = load_sum(dp + (long long)(tt + D) * row, a.ld, A);
        const unsigned ev = ksa_step(c, msgs, t0 + tt + 1, s);
        rp[(long long)tt * a.ld] = s.r;
        ep[(long long)tt * a.ld] = (unsigned char)ev;
      }
    }
    if (next < kChecks && t + D == check_tick(next)) {  // checkpoint ticks are multiples of D
      const TpState mine{s.r, s.su, s.sd};
      if (same_future(mine, rec->ck[next], t0 + t + D)) return next;
      ++next;
      if (next < kChecks && !has_check(len, next)) next = kChecks;
    }
  }
  return -1;
}

// ---- pass 2 ------------------------------------------------------------------------------------------
template <int A>
__global__ void __launch_bounds__(BLOCK) tp_fix_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const int k = blockIdx.y + 1;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const ksa_step_consts c(a.pk);
  TpChunk* const rec = &g_tp_scratch[(long long)k * a.N + n];
  const TpState entry = g_tp_scratch[(long long)(k - 1) * a.N + n].exit1;
  const TpState guess{a.st_r[n], 0, 0};
  if (same_future(entry, guess, t0)) return;  // status stays 0: pass 1 ran this chunk from an equivalent state
  ksa_trace_regs s{entry.r, entry.su, entry.sd};
  const int j = tp_replay_until_match<A>(a, c, n, t0, len, rec, 0, s);
  rec->stop2 = TpState{s.r, s.su, s.sd};
  rec->status = j >= 0 ? j + 1 : -1;
}

// ---- pass 3 ------------------------------------------------------------------------------------------
// A stamp value s was taken by a scale call after local tick t  <=>  s - c > t  (s = call tick + c).
__device__ __forceinline__ int later_stamp(int keep, int cand, int c, int after) {
  return (cand - c > after) ? cand : keep;
}

template <int A>
__global__ void __launch_bounds__(BLOCK) tp_chain_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const ksa_step_consts c(a.pk);
  const TpChunk* r0 = &g_tp_scratch[n];
  TpState cur = r0->exit1;  // chunk 0 ran from the true state
  for (int k = 1; k < a.K; ++k) {
    const int t0 = k * a.L;
    const int len = min(a.L, a.T - t0);
    const TpChunk* rec = &g_tp_scratch[(long long)k * a.N + n];
    const TpState used = g_tp_scratch[(long long)(k - 1) * a.N + n].exit1;  // pass 2's entry state
    const int status = rec->status;
    if (same_future(used, cur, t0)) {
      // pass 2 started from (the equivalent of) the true state: its work stands
      if (status < 0) {  // it rewrote the whole chunk
        cur.r = rec->stop2.r;
        cur.su = later_stamp(cur.su, rec->stop2.su, c.cu, t0);
        cur.sd = later_stamp(cur.sd, rec->stop2.sd, c.cd, t0);
      } else {
        const int tc = status == 0 ? t0 : t0 + check_tick(status - 1);  // pass 1's trajectory is true after tc
        if (status > 0) {
          cur.su = later_stamp(cur.su, rec->stop2.su, c.cu, t0);
          cur.sd = later_stamp(cur.sd, rec->stop2.sd, c.cd, t0);
        }
        cur.r = rec->exit1.r;
        cur.su = later_stamp(cur.su, rec->exit1.su, c.cu, tc);
        cur.sd = later_stamp(cur.sd, rec->exit1.sd, c.cd, tc);
      }
    } else {
      // an earlier chunk never converged: replay this one from the true state. The outputs are pass 2's
      // up to its checkpoint and pass 1's after it, so only checkpoints from there on can end the replay;
      // if pass 2 rewrote the whole chunk, nothing of pass 1 is left in it.
      ksa_trace_regs s{cur.r, cur.su, cur.sd};
      const int first = status < 0 ? kChecks : (status == 0 ? 0 : status - 1);
      const int j = tp_replay_until_match<A>(a, c, n, t0, len, rec, first, s);
      cur.su = s.su;
      cur.sd = s.sd;
      if (j >= 0) {
        const int tc = t0 + check_tick(j);
        cur.r = rec->exit1.r;
        cur.su = later_stamp(cur.su, rec->exit1.su, c.cu, tc);
        cur.sd = later_stamp(cur.sd, rec->exit1.sd, c.cd, tc);
      } else {
        cur.r = s.r;
      }
    }
  }
  a.st_r[n] = cur.r;
  a.st_ju[n] = ksa_global_stamp(cur.su, c.cu, a.tick0, a.st_ju[n]);
  a.st_jd[n] = ksa_global_stamp(cur.sd, c.cd, a.tick0, a.st_jd[n]);
}

// One time-parallel replay per device at a time (the scratch is shared): `busy` is recorded after
// the last launch of a call and queried by the next one.
constexpr int kMaxDevices = 64;
std::mutex g_tp_mutex;
cudaEvent_t g_tp_busy[kMaxDevices] = {};
cudaStream_t g_tp_stream[kMaxDevices] = {};  // stream of the call that recorded `busy`

template <int A>
int tp_launch(const TpArgs& a, cudaStream_t st) {
  const unsigned gx = (unsigned)((a.N + BLOCK - 1) / BLOCK);
  tp_spec_kernel<A><<<dim3(gx, (unsigned)a.K), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_spec_kernel");
  tp_fix_kernel<A><<<dim3(gx, (unsigned)(a.K - 1)), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_fix_kernel");
  tp_chain_kernel<A><<<gx, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_chain_kernel");
  return KSA_OK;
}

}  // namespace

// Number of chunks the time-parallel replay would use for this call, or 0 if the plain kernels should
// run (too many traces to gain, too few ticks, or no room in the scratch). KSA_TIMEPAR_CHUNKS=<K>
// (environment; tests and measurements) forces K chunks whenever that is possible at all, =0 turns
// the path off.
int ksa_timepar_chunks(int64_t T, int64_t N) {
  int forced = -1;
  if (const char* e = std::getenv("KSA_TIMEPAR_CHUNKS")) forced = std::atoi(e);
  if (forced == 0) return 0;
  long long K;
  if (forced > 0) {
    K = forced;
    if (K < 2 || N * K > kMaxChunks || (T + K - 1) / K < 2 * D) return 0;
  } else {
    // one trace per thread fills the GPU at about 2^18 threads (profiles/r01_small_n_variants.txt:
    // 0.82 of the copy peak at 131072 traces, 0.73 at 65536); chunks of at least 1024 ticks keep the
    // re-replayed prefixes (32 ticks in the common case) and the three launches small against the work
    if (N > 98304 || T < 4096) return 0;
    K = 1;
    while (N * K < (1 << 18) && T / (2 * K) >= 1024 && K < 256) K *= 2;
    if (K < 2 || N * K > kMaxChunks) return 0;
  }
  return (int)K;
}

// ksa_replay_full for (T, N) with both outputs and no failure mask, in K chunks per trace. Returns
// KSA_OK, an error, or 1 if the scratch is in use by an earlier call on this device (or the stream is
// being captured): the caller then takes the plain path.
int ksa_timepar_replay_full(const ksa_policy_k& pk, const int64_t* depth_dev, int64_t T, int32_t A, int64_t N,
                            int64_t ld, int64_t tick0, const ksa_state* state, int32_t* replicas_dev,
                            uint8_t* events_dev, int K, cudaStream_t st) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) {
    cudaGetLastError();
    return 1;
  }
  if (cap != cudaStreamCaptureStatusNone) return 1;
  int device = -1;
  KSA_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= kMaxDevices) return 1;
  std::lock_guard<std::mutex> lock(g_tp_mutex);
  if (!g_tp_busy[device]) {
    KSA_CUDA(cudaEventCreateWithFlags(&g_tp_busy[device], cudaEventDisableTiming));
  } else if (cudaEventQuery(g_tp_busy[device]) != cudaSuccess) {
    cudaGetLastError();
    // On the same stream the new call is ordered behind the old one anyway; on another stream the
    // two would overlap on the scratch: plain path.
    if (st != g_tp_stream[device]) return 1;
  }
  TpArgs a;
  a.pk = pk;
  a.depth = (const long long*)depth_dev;
  a.ld = ld;
  a.tick0 = tick0;
  a.N = N;
  a.T = (int)T;
  a.A = A;
  a.K = K;
  a.L = (int)(((T + K - 1) / K + D - 1) / D * D);
  a.one = 1;
  a.st_r = state->replicas_dev;
  a.st_ju = (long long*)state->last_up_tick_dev;
  a.st_jd = (long long*)state->last_down_tick_dev;
  a.replicas = replicas_dev;
  a.events = events_dev;
  if ((long long)a.L * (K - 1) >= T) return 1;  // rounding left the last chunk empty: not worth a special case
  int rc = A == 1 ? tp_launch<1>(a, st) : A == 2 ? tp_launch<2>(a, st) : tp_launch<3>(a, st);
  if (rc != KSA_OK) return rc;
  KSA_CUDA(cudaEventRecord(g_tp_busy[device], st));
  g_tp_stream[device] = st;
  return KSA_OK;
}
