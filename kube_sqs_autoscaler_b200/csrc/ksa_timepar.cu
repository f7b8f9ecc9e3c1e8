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
//  pass 1  thread (8 traces or 1, chunk): start `warm` ticks BEFORE the chunk from a guessed state (the
//          trace's starting replica count, both gates open) and replay up to the chunk without
//          writing anything: most runs have met the true trajectory by then. Record the state at the
//          chunk's first tick, replay the chunk writing its outputs, record the state after a few
//          checkpoint ticks and after the last. Chunk 0 starts from the true state.
//  pass 2a thread (trace, chunk >= 1): the previous chunk's pass-1 exit state is this chunk's entry
//          state. If it has the same future as the state pass 1 entered the chunk with, the chunk is
//          done; otherwise it goes on a list.
//  pass 2b thread (list item): replay from the entry state, rewriting the outputs, until the state at a
//          checkpoint has the same future as what pass 1 left there: from that tick on pass 1's
//          outputs are already right. If no checkpoint matches, the whole chunk is rewritten.
//  pass 3  thread (trace): if every chunk converged (2a or 2b), all entry states were the true ones by
//          induction, and the carry state is the last chunk's exit with the latest TRUE stamps.
//          Otherwise walk the chunks in order with the true state and replay again what was
//          replayed from a wrong entry.
// Results are bit-identical to the sequential replay whatever K and `warm` are and whether or not
// any chunk converges (a trace that never does is simply replayed sequentially by pass 3); the
// tests force small chunks on adversarial policies (one-sided clamps with r0 outside [min, max],
// min > max, cool-downs longer than the call) to check exactly that.
//
// "Same future" at tick t means equal replica counts and equal gates for every later tick: a stamp
// that opens its gate at or before tick t + 1 is as good as any other such stamp. The stamps
// written back to the caller are the true ones: per direction, the stamp of the last scale call of
// the true trajectory (a pass-1 stamp only counts if it was taken after the tick from which pass
// 1's trajectory is the true one; stamps taken during the warm-up never do).
//
// Scratch: one 64-byte record per (trace, chunk) and a 4-byte list slot, in fixed device arrays of
// the library image (kMaxChunks records: 34 MB per GPU), so the entry point still allocates nothing.
// One time-parallel replay per device at a time: a call that finds the scratch busy on another
// stream takes the plain kernels instead.
#include "ksa_internal.h"

#include <cstdlib>
#include <mutex>

namespace {

constexpr int kMaxChunks = 1 << 19;   // records: traces x chunks of one call
constexpr int kChecks = 3;
__host__ __device__ constexpr int check_tick(int j) { return 64 << (2 * j); }  // 64, 256, 1024 ticks into the chunk
constexpr int kMargin = 8;            // a chunk has checkpoint j if check_tick(j) + kMargin <= its length
constexpr int BLOCK = 128;

struct TpState {
  int r, su, sd;  // replicas; first local tick at which the up / down gate is open (ksa_policy.cuh)
};
struct TpChunk {
  int status;        // 0 entry had the future pass 1 assumed; j+1 pass 2b converged at checkpoint j; -1 never
  TpState entry1;    // pass 1: state at the chunk's first tick. Pass 2b reuses the slot for the state it stopped with
  TpState exit1;     // pass 1: state after the chunk's last tick
  TpState ck[kChecks];  // pass 1: state after chunk tick check_tick(j), if the chunk has it
};
static_assert(sizeof(TpChunk) == 64, "scratch record");

static __device__ __align__(256) TpChunk g_tp_scratch[kMaxChunks];
static __device__ int g_tp_list[kMaxChunks];  // (trace, chunk) items pass 2b has to replay
static __device__ int g_tp_count;

struct TpArgs {
  ksa_policy_k pk;
  const long long* depth;
  long long ld, tick0, N;
  int T, A, K, L;  // ticks of the call, attribute planes, chunks, ticks per chunk (a multiple of 8)
  int warm;        // ticks replayed before a chunk without output (a multiple of 8, <= L)
  int one;
  int* st_r;
  long long* st_ju;
  long long* st_jd;
  int* replicas;
  unsigned char* events;
};

// index of the checkpoint reached after `done` ticks of a chunk of `len`, or -1
__device__ __forceinline__ int check_index(int done, int len) {
  if (done + kMargin > len) return -1;
#pragma unroll
  for (int j = 0; j < kChecks; ++j)
    if (done == check_tick(j)) return j;
  return -1;
}

// Same behaviour from local tick t + 1 on.
__device__ __forceinline__ bool same_future(const TpState& a, const TpState& b, int t) {
  return a.r == b.r && max(a.su, t + 1) == max(b.su, t + 1) && max(a.sd, t + 1) == max(b.sd, t + 1);
}

__device__ __forceinline__ long long load_sum(const long long* p, long long ld, int A) {
  unsigned long long acc = (unsigned long long)__ldg(p);  // sqs/sqs.go:56-64, wrapping
  for (int a = 1; a < A; ++a) acc += (unsigned long long)__ldg(p + a * ld);
  return (long long)acc;
}

__device__ __forceinline__ TpChunk* record(const TpArgs& a, int k, long long n) {
  return &g_tp_scratch[(long long)k * a.N + n];
}

// ---- pass 1, V traces per thread (V = 8: 256-bit loads and stores as in the main kernel, needs the
// vector layout; V = 1: any layout) -----------------------------------------------------------------
template <int V>
__device__ __forceinline__ void tp_load(const long long* p, long long (&v)[V]) {
  if constexpr (V == 1) {
    asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(v[0]) : "l"(p));
  } else {
    static_assert(V == 8, "V is 1 or 8");
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[4]), "=l"(v[5]), "=l"(v[6]), "=l"(v[7]) : "l"(p + 4));
  }
}
__device__ __forceinline__ unsigned tp_pack4(unsigned e0, unsigned e1, unsigned e2, unsigned e3) {
  unsigned w;  // multiply-adds: FMA pipe instead of shifts and LOP3 on the ALU pipe
  asm("mad.lo.u32 %0, %2, 256, %1;\n\t"
      "mad.lo.u32 %0, %3, 65536, %0;\n\t"
      "mad.lo.u32 %0, %4, 16777216, %0;"
      : "=&r"(w) : "r"(e0), "r"(e1), "r"(e2), "r"(e3));
  return w;
}
template <int V>
__device__ __forceinline__ void tp_store(int* rp, unsigned char* ep, const int (&r)[V], const unsigned (&e)[V]) {
  if constexpr (V == 1) {
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(rp), "r"(r[0]) : "memory");
    asm volatile("st.global.cs.b8 [%0], %1;" ::"l"(ep), "r"(e[0]) : "memory");
  } else {
    asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(rp), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
    const unsigned w0 = tp_pack4(e[0], e[1], e[2], e[3]), w1 = tp_pack4(e[4], e[5], e[6], e[7]);
    asm volatile("st.global.cs.v2.b32 [%0], {%1,%2};" ::"l"(ep), "r"(w0), "r"(w1) : "memory");
  }
}

template <int A, int V, int D, int TPB, int MINB>
__global__ void __launch_bounds__(TPB, MINB) tp_spec_kernel(const __grid_constant__ TpArgs a) {
  const long long n0 = ((long long)blockIdx.x * TPB + threadIdx.x) * V;
  if (n0 >= a.N) return;  // (V = 8: N is a multiple of 8)
  const int k = blockIdx.y;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const int tb = k == 0 ? 0 : t0 - a.warm;  // first tick replayed
  const int te = t0 + len;
  const ksa_step_consts c(a.pk);
  const int one = a.one;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_tp_count = 0;  // pass 2a's list starts empty
  ksa_trace_regs s[V];
#pragma unroll
  for (int v = 0; v < V; ++v) {
    s[v].r = a.st_r[n0 + v];
    if (k == 0) {
      s[v].su = ksa_local_stamp(a.pk.cu, c.cu, a.tick0, a.st_ju[n0 + v]);
      s[v].sd = ksa_local_stamp(a.pk.cd, c.cd, a.tick0, a.st_jd[n0 + v]);
    } else {
      s[v].su = s[v].sd = 0;  // the guess: the trace's starting replica count, both gates open
    }
  }
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + n0;
  int* rp = a.replicas + n0;
  unsigned char* ep = a.events + n0;

  // Register ring over the ticks tb, tb + 1, ...: slot (tt - tb) % D.
  long long buf[D][A][V];
#pragma unroll
  for (int j = 0; j < D; ++j)
    if (tb + j < te) {
#pragma unroll
      for (int p = 0; p < A; ++p) tp_load<V>(dp + (long long)(tb + j) * row + p * a.ld, buf[j][p]);
    }
  auto tick = [&](const int tt, const int j, const bool prefetch, const bool store) {
    long long msgs[V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      unsigned long long acc = (unsigned long long)buf[j][0][v];  // sqs.go:56-64, wrapping
#pragma unroll
      for (int p = 1; p < A; ++p) acc += (unsigned long long)buf[j][p][v];
      msgs[v] = (long long)acc;
    }
    if (prefetch) {
#pragma unroll
      for (int p = 0; p < A; ++p) tp_load<V>(dp + (long long)(tt + D) * row + p * a.ld, buf[j][p]);
    }
    int rr[V];
    unsigned ev[V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      ev[v] = ksa_step_fma_byte(c, msgs[v], tt + 1, one, s[v]);
      rr[v] = s[v].r;
    }
    if (store) tp_store<V>(rp + (long long)tt * a.ld, ep + (long long)tt * a.ld, rr, ev);
  };
  int t = tb;
  // warm-up: no output (warm is a multiple of D; the last chunk may be shorter than the ring)
#pragma unroll 1
  for (; t < t0; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) tick(t + j, j, t + j + D < te, false);
  }
#pragma unroll
  for (int v = 0; v < V; ++v) record(a, k, n0 + v)->entry1 = TpState{s[v].r, s[v].su, s[v].sd};
  // the chunk: steady state, then the drain with guards
#pragma unroll 1
  for (; t + 2 * D <= te; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) tick(t + j, j, true, true);
    const int q = check_index(t + D - t0, len);
    if (q >= 0) {
#pragma unroll
      for (int v = 0; v < V; ++v) record(a, k, n0 + v)->ck[q] = TpState{s[v].r, s[v].su, s[v].sd};
    }
  }
#pragma unroll 1
  for (; t < te; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      if (t + j < te) tick(t + j, j, t + j + D < te, true);
    const int q = check_index(t + D - t0, len);  // (a checkpoint is never inside the last kMargin ticks)
    if (q >= 0) {
#pragma unroll
      for (int v = 0; v < V; ++v) record(a, k, n0 + v)->ck[q] = TpState{s[v].r, s[v].su, s[v].sd};
    }
  }
#pragma unroll
  for (int v = 0; v < V; ++v) {
    record(a, k, n0 + v)->exit1 = TpState{s[v].r, s[v].su, s[v].sd};
    record(a, k, n0 + v)->status = 0;
  }
}

// Replay chunk (t0, len) from state s, rewriting the outputs, in groups of 8 ticks with the depths
// of the next group already in flight (a plain loop waits a memory latency per tick). Stops after a
// checkpoint tick j >= first_check whose pass-1 state has the same future; returns that j, or -1
// when the chunk ends first.
template <int A>
__device__ int tp_replay_until_match(const TpArgs& a, const ksa_step_consts& c, long long n, int t0, int len,
                                     const TpChunk* rec, int first_check, ksa_trace_regs& s) {
  constexpr int G = 8;
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + (long long)t0 * row + n;
  int* rp = a.replicas + (long long)t0 * a.ld + n;
  unsigned char* ep = a.events + (long long)t0 * a.ld + n;
  long long buf[G];
#pragma unroll
  for (int j = 0; j < G; ++j) buf[j] = j < len ? load_sum(dp + (long long)j * row, a.ld, A) : 0;
  for (int t = 0; t < len; t += G) {
#pragma unroll
    for (int j = 0; j < G; ++j) {
      const int tt = t + j;
      if (tt < len) {
        const long long msgs = buf[j];
        if (tt + G < len) buf[j] = load_sum(dp + (long long)(tt + G) * row, a.ld, A);
        const unsigned ev = ksa_step(c, msgs, t0 + tt + 1, s);
        rp[(long long)tt * a.ld] = s.r;
        ep[(long long)tt * a.ld] = (unsigned char)ev;
      }
    }
    const int q = check_index(t + G, len);  // checkpoint ticks are multiples of 8
    if (q >= first_check && q >= 0) {
      const TpState mine{s.r, s.su, s.sd};
      if (same_future(mine, rec->ck[q], t0 + t + G)) return q;
    }
  }
  return -1;
}

// ---- pass 2a -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) tp_check_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  const int k = blockIdx.y + 1;
  bool need = false;
  if (n < a.N) {
    const TpState entry = record(a, k - 1, n)->exit1;
    need = !same_future(entry, record(a, k, n)->entry1, k * a.L);
  }
  // one atomic per warp
  const unsigned vote = __ballot_sync(0xffffffffu, need);
  if (vote == 0u) return;
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == __ffs(vote) - 1) base = atomicAdd(&g_tp_count, __popc(vote));
  base = __shfl_sync(0xffffffffu, base, __ffs(vote) - 1);
  if (need) g_tp_list[base + __popc(vote & ((1u << lane) - 1u))] = (int)((long long)k * a.N + n);
}

// ---- pass 2b -------------------------------------------------------------------------------------
template <int A>
__global__ void __launch_bounds__(BLOCK) tp_fix_kernel(const __grid_constant__ TpArgs a) {
  const int idx = blockIdx.x * BLOCK + threadIdx.x;
  if (idx >= g_tp_count) return;
  const int item = g_tp_list[idx];
  const int k = (int)(item / a.N);
  const long long n = item % a.N;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const ksa_step_consts c(a.pk);
  TpChunk* const rec = record(a, k, n);
  const TpState entry = record(a, k - 1, n)->exit1;
  ksa_trace_regs s{entry.r, entry.su, entry.sd};
  const int j = tp_replay_until_match<A>(a, c, n, t0, len, rec, 0, s);
  rec->entry1 = TpState{s.r, s.su, s.sd};  // (the slot is free: pass 2a has read it)
  rec->status = j >= 0 ? j + 1 : -1;
}

// ---- pass 3 --------------------------------------------------------------------------------------
// A stamp value s was taken by a scale call after local tick t  <=>  s - c > t  (s = call tick + c).
__device__ __forceinline__ int later_stamp(int keep, int cand, int c, int after) {
  return (cand - c > after) ? cand : keep;
}

template <int A>
__global__ void __launch_bounds__(BLOCK) tp_chain_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const ksa_step_consts c(a.pk);
  TpState cur = record(a, 0, n)->exit1;  // chunk 0 ran from the true state
  // Every chunk converged (the common case): all entry states were the true ones, the replica count is
  // the last chunk's, and each stamp is the latest one taken where the trajectory was the true one.
  // Valid candidates are later than anything before them, so "latest" is a maximum and the loop has
  // no dependence from one chunk's record to the next one's loads.
  bool all = true;
  int su = cur.su, sd = cur.sd;
  for (int k = 1; k < a.K; ++k) {
    const TpChunk* rec = record(a, k, n);
    const int status = rec->status;
    const int t0 = k * a.L;
    all = all && status >= 0;
    const int tc = status <= 0 ? t0 : t0 + check_tick(status - 1);
    if (status > 0) {  // pass 2b's own replay, from the chunk's first tick
      su = max(su, later_stamp(INT_MIN, rec->entry1.su, c.cu, t0));
      sd = max(sd, later_stamp(INT_MIN, rec->entry1.sd, c.cd, t0));
    }
    su = max(su, later_stamp(INT_MIN, rec->exit1.su, c.cu, tc));
    sd = max(sd, later_stamp(INT_MIN, rec->exit1.sd, c.cd, tc));
  }
  if (all) {
    cur.r = record(a, a.K - 1, n)->exit1.r;
    cur.su = su;
    cur.sd = sd;
  } else {
    // Some chunk never converged: walk the chunks in order with the true state.
    for (int k = 1; k < a.K; ++k) {
      const int t0 = k * a.L;
      const int len = min(a.L, a.T - t0);
      const TpChunk* rec = record(a, k, n);
      const TpState used = record(a, k - 1, n)->exit1;  // pass 2's entry state
      const int status = rec->status;
      if (same_future(used, cur, t0)) {
        // pass 2 started from (the equivalent of) the true state: its work stands
        if (status < 0) {  // 2b rewrote the whole chunk; its end state is in the entry1 slot
          cur.r = rec->entry1.r;
          cur.su = later_stamp(cur.su, rec->entry1.su, c.cu, t0);
          cur.sd = later_stamp(cur.sd, rec->entry1.sd, c.cd, t0);
        } else {
          const int tc = status == 0 ? t0 : t0 + check_tick(status - 1);  // pass 1's trajectory is true after tc
          if (status > 0) {
            cur.su = later_stamp(cur.su, rec->entry1.su, c.cu, t0);
            cur.sd = later_stamp(cur.sd, rec->entry1.sd, c.cd, t0);
          }
          cur.r = rec->exit1.r;
          cur.su = later_stamp(cur.su, rec->exit1.su, c.cu, tc);
          cur.sd = later_stamp(cur.sd, rec->exit1.sd, c.cd, tc);
        }
      } else {
        // an earlier chunk never converged: replay this one from the true state. The outputs are pass 2b's
        // up to its checkpoint and pass 1's after it, so only checkpoints from there on can end the replay;
        // if pass 2b rewrote the whole chunk, nothing of pass 1 is left in it.
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
int tp_launch(const TpArgs& a, bool wide, cudaStream_t st) {
  if (wide) {  // 8 traces per thread, the launch shape of the main kernel's default
    constexpr int TPB = 256;
    const unsigned gx = (unsigned)((a.N / 8 + TPB - 1) / TPB);
    tp_spec_kernel<A, 8, A == 1 ? 2 : 1, TPB, A == 1 ? 2 : 1><<<dim3(gx, (unsigned)a.K), TPB, 0, st>>>(a);
  } else {
    const unsigned gx = (unsigned)((a.N + BLOCK - 1) / BLOCK);
    tp_spec_kernel<A, 1, 8, BLOCK, A == 1 ? 8 : (A == 2 ? 5 : 4)><<<dim3(gx, (unsigned)a.K), BLOCK, 0, st>>>(a);
  }
  KSA_LAUNCH_CHECK("tp_spec_kernel");
  const unsigned gn = (unsigned)((a.N + BLOCK - 1) / BLOCK);
  tp_check_kernel<<<dim3(gn, (unsigned)(a.K - 1)), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_check_kernel");
  const long long items = a.N * (a.K - 1);  // (the list's length is only known on the device: idle threads return at once)
  tp_fix_kernel<A><<<(unsigned)((items + BLOCK - 1) / BLOCK), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_fix_kernel");
  tp_chain_kernel<A><<<gn, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_chain_kernel");
  return KSA_OK;
}

}  // namespace

// How the time-parallel replay would cut this call: chunks K (0 = take the plain kernels: too many
// traces to gain, too few ticks, or no room in the scratch), ticks per chunk L, warm-up ticks, and
// whether pass 1 can run 8 traces per thread (`vector_ok`: the layout the vector kernels need and N
// a multiple of 8). Environment, for tests and measurements: KSA_TIMEPAR_CHUNKS=<K> forces K chunks
// whenever that is possible at all (0 turns the path off), KSA_TIMEPAR_WARM=<ticks> the warm-up,
// KSA_TIMEPAR_WIDE=0 one trace per thread.
int ksa_timepar_plan(int64_t T, int64_t N, bool vector_ok, int* L_out, int* warm_out, int* wide_out) {
  int forced = -1;
  if (const char* e = std::getenv("KSA_TIMEPAR_CHUNKS")) forced = std::atoi(e);
  if (forced == 0) return 0;
  bool wide = vector_ok;
  if (const char* e = std::getenv("KSA_TIMEPAR_WIDE")) wide = wide && std::atoi(e) != 0;
  int warm = 128;  // profiles/r02_small_n.txt: 0.80 of the copy peak at 4096 x 10^5 against 0.78 with 64, 0.68 with none
  if (const char* e = std::getenv("KSA_TIMEPAR_WARM")) warm = std::max(0, std::atoi(e)) / 8 * 8;
  long long K;
  if (forced > 0) {
    K = forced;
    if (K < 2 || (T + K - 1) / K < 16) return 0;
  } else {
    // One trace per thread fills the GPU at about 2^18 threads, eight per thread (1-KB requests, the
    // shape that reaches the copy peak) at about 400k trace-chunks (profiles/r01_small_n_variants.txt).
    // Chunks of at least 8 x warm ticks keep the warm-up and the re-replayed prefixes small.
    if (N > 196608 || T < 4096) return 0;
    const long long want = wide ? 393216 : 262144;
    K = (want + N - 1) / N;
    const long long kmax = T / (8 * std::max(warm, 64));
    if (K > kmax) K = kmax;
    if (K > 512) K = 512;
    if (K < 2) return 0;
  }
  long long L = ((T + K - 1) / K + 7) / 8 * 8;
  K = (T + L - 1) / L;  // (rounding L up may leave fewer chunks)
  if (K < 2 || N * K > kMaxChunks) return 0;
  if (warm > L) warm = (int)L;
  *L_out = (int)L;
  *warm_out = warm;
  *wide_out = wide ? 1 : 0;
  return (int)K;
}

// ksa_replay_full for (T, N) with both outputs and no failure mask, in K chunks per trace. Returns
// KSA_OK, an error, or 1 if the scratch is in use by an earlier call on another stream of this
// device (or the stream is being captured): the caller then takes the plain path.
int ksa_timepar_replay_full(const ksa_policy_k& pk, const int64_t* depth_dev, int64_t T, int32_t A, int64_t N,
                            int64_t ld, int64_t tick0, const ksa_state* state, int32_t* replicas_dev,
                            uint8_t* events_dev, int K, int L, int warm, int wide, cudaStream_t st) {
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
  a.L = L;
  a.warm = warm;
  a.one = 1;
  a.st_r = state->replicas_dev;
  a.st_ju = (long long*)state->last_up_tick_dev;
  a.st_jd = (long long*)state->last_down_tick_dev;
  a.replicas = replicas_dev;
  a.events = events_dev;
  int rc = A == 1 ? tp_launch<1>(a, wide != 0, st) : A == 2 ? tp_launch<2>(a, wide != 0, st) : tp_launch<3>(a, wide != 0, st);
  if (rc != KSA_OK) return rc;
  KSA_CUDA(cudaEventRecord(g_tp_busy[device], st));
  g_tp_stream[device] = st;
  return KSA_OK;
}
