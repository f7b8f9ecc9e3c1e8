// ksa_timepar.cu — replay for trace counts that cannot fill the GPU: the time axis is cut into K
// chunks per trace and the chunks are replayed in parallel. First the full-trajectory replay.
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
//          outputs are already right. If no checkpoint matches, the whole chunk is rewritten. Only
//          when few chunks are listed: many mean that the guesses fail for a reason.
//  pass 3a thread (trace), for traces with chunks that neither stand nor got patched: walk the chunks in
//          order with the true state, O(1) per chunk. A chunk replayed from the right gates but a
//          replica count that is off, and without a scale call, is shifted by the offset (the trace
//          that goes quiet: its state never forgets). Any other chunk that does not fit is marked to
//          be replayed a second time from the true state; the state after it is worked out for a chunk
//          in which every tick met the same threshold (the saturated and the idle queue: calls exactly
//          one cool-down apart, a guess keeps a wrong phase for ever; pass 1 notes such chunks, pure_exit),
//          assumed to be what the chunk ended with the first time otherwise.
//  then    the second replay: pass 1's kernel, one trace per thread, over the marked chunks; and, for
//          marked chunks in which every tick met the same threshold, no replay at all: their outputs
//          have a closed form and are written without reading a depth (tp_fill_kernel: 8 adjacent
//          traces per thread, 256-bit stores, where a whole group has the form).
//  pass 3b thread (trace): if every chunk stands or was patched, all entry states were the true ones by
//          induction, and the carry state is the last chunk's exit with the latest TRUE stamps.
//          Otherwise the same walk over what is there now: wrong assumptions of pass 3a show as chunks
//          that still do not fit; up to two are replayed here, a trace that needs more is flagged
//          and left alone.
//  then    the replica counts of the chunks marked for an offset: a chunk without a scale call has ONE
//          count from its first tick to its last, which is written where 8 adjacent traces have such
//          a chunk (tp_shift_fill_kernel) and read, added to and written back elsewhere (tp_shift_kernel);
//          and the one-trace-per-thread kernel of ksa_replay.cu over the flagged traces only, from
//          the caller's state. (Before passes 3a / second replay, a single trace that never met its
//          true trajectory made the call several times SLOWER than the plain kernels: pass 3 replayed
//          its 10^5 ticks in one thread.)
//  first   above 16384 traces, where two chunked replays cost more than one plain one: a probe
//          (tp_probe_kernel) of whether guesses meet the truth on these depths at all; if not, every
//          pass returns at once and the plain kernel does the call.
// Results are bit-identical to the sequential replay whatever K and `warm` are and whether or not
// any chunk converges; the tests force small chunks on adversarial policies (one-sided clamps with
// r0 outside [min, max], min > max, cool-downs longer than the call, traces that go quiet, saturated
// and idle queues, random walks) to check exactly that.
//
// "Same future" at tick t means equal replica counts and equal gates for every later tick: a stamp
// that opens its gate at or before tick t + 1 is as good as any other such stamp. The stamps
// written back to the caller are the true ones: per direction, the stamp of the last scale call of
// the true trajectory (a pass-1 stamp only counts if it was taken after the tick from which pass
// 1's trajectory is the true one; stamps taken during the warm-up never do).
//
// Scratch: per (trace, chunk) a 64-byte record, a list slot and a few words, in fixed device arrays of
// the library image (kMaxChunks slots: 44 MB per GPU; summary mode, further down, adds as many
// ksa_summary records: 76 MB), so the entry points still allocate nothing.
// One time-parallel replay per device at a time: a call that finds the scratch busy on another
// stream takes the plain kernels instead.
//
// The same passes serve ksa_replay_full_faults (FAULTS: the failure mask is read next to the
// depths; a failed call changes no state, so a run from a wrong state still meets the true one).
// ksa_replay_summary / ksa_replay_summary_faults have no trajectory to patch and are built
// differently: second half of this file.
#include "ksa_internal.h"

#include <cstdlib>
#include <mutex>

namespace {

constexpr int kMaxChunks = 1 << 19;   // records: traces x chunks of one call
constexpr int kChecks = 3;
__host__ __device__ constexpr int check_tick(int j) { return 64 << (2 * j); }  // 64, 256, 1024 ticks into the chunk
constexpr int kMargin = 8;            // a chunk has checkpoint j if check_tick(j) + kMargin <= its length
constexpr int BLOCK = 128;
// Trace counts up to which the chunked path is taken by default: where it stops paying against one
// trace per thread over the whole call (profiles/r02_small_n_modes.txt).
// (summary mode: on depths with long stretches on one side of a threshold many chunks are replayed twice;
// the limit is where that still costs no more than the plain kernel: random-walk depths, 32768 traces:
// 1.76 ms against 1.69; at 49152: 2.00 against 1.49, while independent depths gain 0.66 -> 0.80 there)
constexpr long long kTpMaxSummary = 32768;
// With a failure mask there is no closed form for saturated / idle stretches (a failed call takes no stamp and
// moves the phase), and runs from different phases meet only when failures happen to line them up: at 32768
// traces of such queues the chunked passes cost 3.6 ms against 1.9 (profiles/r02_small_n_robust.txt)
constexpr long long kTpMaxSummaryFaults = 16384;
// Full mode, with or without a mask: the limit is set by the depths on which chunks are replayed twice or
// shifted, not by independent ones (profiles/r02_small_n_robust.txt, plain -> chunked ms; independent /
// random walk / quiet / saturated-idle): 32768 traces 2.35 -> 1.68 / 2.45 / 2.10 / 2.45; 65536: 1.87 -> 1.67 /
// 1.96 / 2.04 / 1.97; 98304: 2.24 -> 2.45 / 2.34 / 2.84 / 2.34 (the plain kernels are at 0.88 of the copy peak
// there). With a mask, 49152: 2.73 -> 2.31 / 2.74 / 2.64 / 2.85; 65536: 2.04 -> 2.01 / 2.44 / 2.22 / 2.16.
constexpr long long kTpMaxFull = 49152;
constexpr long long kTpMaxFullFaults = 49152;
// ... and again where one trace per thread starts a second, nearly empty wave of 148 x 1024 threads
// (163840 traces: 4.31 -> 2.91 / 4.35 / 3.51 / 4.36; with a mask 3.96 -> 3.69 / 4.10 / 3.99 / 4.05), as long as
// two chunks still fit one wave of pass 1
constexpr long long kTpFullSecondWave = 151552;
constexpr long long kTpFullFaultsSecondWave = 151552;

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
// Pass 3's verdict per trace: 0 = the call's results are written; 1 = this trace never met its true
// trajectory within the few chunks pass 3 may replay itself: nothing was written for it, and the
// one-trace-per-thread kernel that follows the passes (launched for flagged traces only) replays it
// from the caller's state. Without that, one such trace made the whole call several times slower
// than the plain kernels (pass 3 replaying 10^5 ticks in one thread: profiles/r02_small_n_modes.txt).
constexpr int kMaxTraces = 196608;
static __device__ int g_tp_unresolved[kMaxTraces];
constexpr int kChainReplays = 2;  // chunks pass 3 replays by itself per trace before it gives the trace up

// The offset pass 3 decided to shift a chunk's replica counts by (0: none; pass 2a clears it), applied by
// tp_shift_kernel.
static __device__ int g_tp_delta[kMaxChunks];
constexpr int kLeftToShift = 100;  // TpChunk::status: pass 2a left the chunk to pass 3, which shifts it
constexpr int kListed = 101;       // listed for pass 2b, which replaces it unless it leaves the chunk alone
constexpr int kRerun = 102;        // replayed a second time from the entry state pass 3a put into entry1
constexpr int kRedone = 104;       // replayed by pass 3b itself, from the true state: entry1 / exit1 are stale
constexpr int kFill = 103;         // every tick met the same threshold: outputs written from the closed form (tp_fill_chunk),
                                   // from the entry state in entry1; exit1 = the state after it, worked out by pass 3a

// What pass 2b replayed a listed chunk from (the previous chunk's exit state at that time: the second
// replay may change that one later), and what kind of chunk it is (noted by pass 1).
static __device__ TpState g_tp_used[kMaxChunks];
static __device__ unsigned char g_tp_kind[kMaxChunks];
// The probe's verdict (tp_probe_kernel): pairs (sampled trace, chunk start) at which warm-ups of three
// lengths did not meet. Above TpArgs::probe_limit every pass returns at once and the plain kernel does the call.
static __device__ int g_tp_probe_bad;
// The trace has a chunk that neither stands as pass 1 wrote it nor was patched up to a checkpoint by pass 2b:
// passes 3a / 3b walk its chunks in order (set by passes 2a / 2b, cleared by pass 3b).
static __device__ int g_tp_walk[kMaxTraces];

struct TpArgs {
  ksa_policy_k pk;
  const long long* depth;
  const unsigned char* fail;  // [T][ld] KSA_FAIL_* bits (the *_faults entry points), else NULL
  long long ld, tick0, N;
  int T, A, K, L;  // ticks of the call, attribute planes, chunks, ticks per chunk (a multiple of 8)
  int warm;        // ticks replayed before a chunk without output (a multiple of 8, <= L)
  int fix_limit;   // pass 2b leaves everything to pass 3a when more chunks than this are listed
  int probe_limit; // INT_MAX: no probe was run
  int probe_traces, probe_stride;  // the probe's sample: groups of 32 adjacent traces, `stride` traces from group to group
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
__device__ __forceinline__ bool same_gates(const TpState& a, const TpState& b, int t) {
  return max(a.su, t + 1) == max(b.su, t + 1) && max(a.sd, t + 1) == max(b.sd, t + 1);
}
__device__ __forceinline__ bool same_future(const TpState& a, const TpState& b, int t) {
  return a.r == b.r && same_gates(a, b, t);
}

// A chunk replayed from the right gates but a replica count that was off by an offset need not be
// replayed again if no scale call happened in it: which calls happen does not depend on the replica
// count (a call that finds the limit reached stamps like any other: main.go:62,76, scale.go:62-65), so
// the true run has none either, and its count is the replayed one plus the offset at every tick, with
// the same events. This is for the trace that goes quiet after some activity — no call for thousands
// of ticks: its state never forgets, so no warm-up makes a guess meet it, and before this test one such
// trace made a call several times slower than the plain kernels (profiles/r02_small_n_modes.txt).
// (tp_shift_safe, summary mode, has the general condition.)
// No scale call in a chunk of pass 1 <=> it left both stamps as they were at the chunk's first tick (a call
// at tick i stamps i + cool-down, later than anything the chunk could have entered with).
__device__ __forceinline__ bool shift_safe(const TpChunk& rec) {
  return rec.exit1.su == rec.entry1.su && rec.exit1.sd == rec.entry1.sd;
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
  } else if constexpr (V == 4) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
  } else {
    static_assert(V == 8, "V is 1, 4 or 8");
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(v[4]), "=l"(v[5]), "=l"(v[6]), "=l"(v[7]) : "l"(p + 4));
  }
}
// V failure-mask bytes of one tick (V = 4: one packed word, V = 8: two)
template <int V>
__device__ __forceinline__ void tp_load_fail(const unsigned char* p, unsigned (&w)[V == 8 ? 2 : 1]) {
  if constexpr (V == 1) {
    asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(w[0]) : "l"(p));
  } else if constexpr (V == 4) {
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w[0]) : "l"(p));
  } else {
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p));
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
  } else if constexpr (V == 4) {
    asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(rp), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3])
                 : "memory");
    const unsigned w0 = tp_pack4(e[0], e[1], e[2], e[3]);
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(ep), "r"(w0) : "memory");
  } else {
    asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(rp), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
    const unsigned w0 = tp_pack4(e[0], e[1], e[2], e[3]), w1 = tp_pack4(e[4], e[5], e[6], e[7]);
    asm volatile("st.global.cs.v2.b32 [%0], {%1,%2};" ::"l"(ep), "r"(w0), "r"(w1) : "memory");
  }
}

// The closed forms below add steps in 64 bits; the reference adds in int32 and wraps (scale.go:66, 94) before it
// clamps. They agree as long as a step from below the limit cannot pass INT_MAX (from above it: INT_MIN),
// which only limits next to the ends of int32 can make happen (found by tools/fuzz_timepar.py: max_pods
// = 2^31 - 1). Otherwise the chunk is replayed.
__device__ __forceinline__ bool step_cannot_wrap(const ksa_step_consts& c, bool up) {
  return up ? (long long)c.max32 + c.up32 <= INT_MAX : (long long)c.min32 - c.down32 >= INT_MIN;
}

// Whether chunks of that kind have the closed forms at all with this policy (pure_exit, tp_fill_chunk).
__device__ __forceinline__ bool pure_steps_ok(const ksa_step_consts& c, int kind) {
  const bool up = kind == 1;
  const int step = up ? c.up32 : c.down32;
  return step >= 0 && step <= (1 << 30) && step_cannot_wrap(c, up);
}

// The outputs of a chunk in which every tick met the scale-up (kind 1) or the scale-down threshold (2),
// from its true entry state, without reading a depth: the cool-down line until the first call (the first
// tick the gate is open) and between calls, one call every cool-down period; a call steps the count
// towards the limit or finds it reached (scale.go:62-69, 89-97).
__device__ void tp_fill_chunk(const TpArgs& a, const ksa_step_consts& c, long long n, int t0, int len, int kind,
                              const TpState& entry) {
  const bool up = kind == 1;
  const int step = up ? c.up32 : c.down32, limit = up ? c.max32 : c.min32;
  const int period = max(up ? c.cu : c.cd, 1);
  const unsigned ev_cool = up ? KSA_EV_UP_COOL : KSA_EV_DOWN_COOL, ev_applied = up ? KSA_EV_UP_APPLIED : KSA_EV_DOWN_APPLIED,
                 ev_noop = up ? KSA_EV_UP_NOOP : KSA_EV_DOWN_NOOP;
  long long next = max(t0 + 1, up ? entry.su : entry.sd);  // local tick of the next call
  int r = entry.r;
  int* rp = a.replicas + (long long)t0 * a.ld + n;
  unsigned char* ep = a.events + (long long)t0 * a.ld + n;
  for (int t = 0; t < len; ++t) {
    unsigned ev = ev_cool;
    if (t0 + t + 1 == next) {
      const bool applied = up ? r < limit : r > limit;
      if (applied) r = up ? (int)min((long long)r + step, (long long)limit) : (int)max((long long)r - step, (long long)limit);
      ev = applied ? ev_applied : ev_noop;
      next += period;
    }
    rp[(long long)t * a.ld] = r;
    ep[(long long)t * a.ld] = (unsigned char)ev;
  }
}

// Can the chunk's outputs be written from the closed form? Marked kFill by pass 3a, or taken by it as pass 1
// replayed it (status 0; kListed / kLeftToShift: pass 2a had doubts that the true state did not confirm) and of a
// kind that has the form: written again from its entry state it comes out as it is (before tp_shift_kernel:
// a chunk that is off by a replica offset is written with that offset).
__device__ __forceinline__ bool tp_fillable(const ksa_step_consts& c, int status, int kind) {
  return status == kFill ||
         ((status == 0 || status == kListed || status == kLeftToShift) && kind != 0 && pure_steps_ok(c, kind));
}
// The two kernels after pass 3a share the marked chunks out by the same rule. A group of 32 adjacent traces
// (a warp of the second replay) with a chunk to be replayed is replayed by that warp, its kFill chunks with
// it. Otherwise a group of 8 adjacent traces whose chunks are all fillable is written by tp_fill_kernel,
// 256-bit stores (a warp: 256 traces, a kilobyte of replica counts per store; the steady state, every queue
// saturated or idle, is all such groups), and what is left over by the second replay's threads one trace
// each (the same bytes written that way everywhere took 0.88 ms at 4096 x 10^5: 2.3 TB/s).
__device__ __forceinline__ bool tp_vector_layout(const TpArgs& a) {
  return a.N % 8 == 0 && a.ld % 8 == 0 && ((unsigned long long)a.replicas & 31u) == 0u && ((unsigned long long)a.events & 7u) == 0u;
}

// Thread (8 adjacent traces, chunk, part of the chunk): the closed form gives the state at any tick, so a
// chunk is cut into kFillParts runs of ticks for as many times the threads (8 traces per thread leave 1168
// warps at 4096 x 10^5, and a warp needs some 100 instructions per tick of its 256 traces: 0.70 ms; in 4
// parts 0.46, in 8 0.43 = 4.7 TB/s of stores).
constexpr int kFillParts = 8;
__device__ __forceinline__ bool pure_exit(const ksa_step_consts& c, int kind, int t0, int len, TpState& g);
__global__ void __launch_bounds__(BLOCK) tp_fill_kernel(const __grid_constant__ TpArgs a) {
  const long long n0 = ((long long)blockIdx.x * BLOCK + threadIdx.x) * 8;
  if (g_tp_probe_bad > a.probe_limit || !tp_vector_layout(a)) return;
  const int k = blockIdx.y + 1;
  const ksa_step_consts c(a.pk);
  int walked = 0;  // (only a trace pass 3a walked has marked chunks: the usual call stops here)
#pragma unroll
  for (int v = 0; v < 8; ++v) walked |= n0 + v < a.N ? g_tp_walk[n0 + v] : 0;
  if (__ballot_sync(0xffffffffu, walked != 0) == 0u) return;
  int kind[8];
  bool rerun = false, any = false, all = true;
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    const bool in = n0 + v < a.N;
    const int status = in ? record(a, k, n0 + v)->status : -1;
    kind[v] = in ? g_tp_kind[(long long)k * a.N + n0 + v] : 0;
    rerun = rerun || status == kRerun;
    any = any || status == kFill;
    all = all && tp_fillable(c, status, kind[v]);
  }
  // lanes 4q .. 4q + 3 hold the 32 traces of one warp of the second replay
  int group_rerun = rerun ? 1 : 0;  // (every lane takes part in both exchanges)
  group_rerun |= __shfl_xor_sync(0xffffffffu, group_rerun, 1);
  group_rerun |= __shfl_xor_sync(0xffffffffu, group_rerun, 2);
  if (group_rerun != 0 || !any || !all) return;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const int part = (len + kFillParts - 1) / kFillParts;
  const int ta = min((int)blockIdx.z * part, len), tb = min(ta + part, len);  // my ticks of the chunk: t0 + ta + 1 .. t0 + tb
  if (ta >= tb) return;
  // per trace: ticks until its next call, its count, and the constants of its direction
  int wait[8], r[8], step[8], limit[8], period[8];
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    const bool up = kind[v] == 1;
    TpState g = record(a, k, n0 + v)->entry1;
    if (ta > 0) pure_exit(c, kind[v], t0, ta, g);  // the state after the ticks before mine
    const long long next = max(t0 + ta + 1, up ? g.su : g.sd);
    wait[v] = (int)min(next - (t0 + ta + 1), (long long)INT_MAX);
    r[v] = g.r;
    step[v] = up ? c.up32 : -c.down32;  // (pure_steps_ok: 0 .. 2^30, and r + step cannot wrap on the side of the limit)
    limit[v] = up ? c.max32 : c.min32;
    period[v] = max(up ? c.cu : c.cd, 1) - 1;
  }
  int* rp = a.replicas + (long long)(t0 + ta) * a.ld + n0;
  unsigned char* ep = a.events + (long long)(t0 + ta) * a.ld + n0;
  for (int t = ta; t < tb; ++t) {
    unsigned ev[8];
#pragma unroll
    for (int v = 0; v < 8; ++v) {
      const bool up = kind[v] == 1;
      const bool call = wait[v] == 0;
      const bool applied = call && (up ? r[v] < limit[v] : r[v] > limit[v]);
      const int stepped = (int)((unsigned)r[v] + (unsigned)step[v]);  // (used only where it cannot wrap)
      const int moved = up ? min(stepped, limit[v]) : max(stepped, limit[v]);
      r[v] = applied ? moved : r[v];
      ev[v] = up ? (call ? (applied ? KSA_EV_UP_APPLIED : KSA_EV_UP_NOOP) : KSA_EV_UP_COOL)
                 : (call ? (applied ? KSA_EV_DOWN_APPLIED : KSA_EV_DOWN_NOOP) : KSA_EV_DOWN_COOL);
      wait[v] = call ? period[v] : wait[v] - 1;
    }
    tp_store<8>(rp, ep, r, ev);
    rp += a.ld;
    ep += a.ld;
  }
}

// RERUN: the second replay. Only the chunks pass 3a marked (kRerun), from the entry state it put into
// entry1, no warm-up; one trace per thread. Chunks marked kFill are written from the closed form instead
// (here or by tp_fill_kernel, see above), unless a chunk of the same warp is replayed: then they are replayed with it.
template <int A, int V, int D, int TPB, int MINB, bool FAULTS, bool RERUN = false>
__global__ void __launch_bounds__(TPB, MINB) tp_spec_kernel(const __grid_constant__ TpArgs a) {
  const long long n0 = ((long long)blockIdx.x * TPB + threadIdx.x) * V;
  if (n0 >= a.N || g_tp_probe_bad > a.probe_limit) return;  // (V > 1: N is a multiple of 8)
  const int k = RERUN ? blockIdx.y + 1 : blockIdx.y;
  if constexpr (RERUN) {
    static_assert(V == 1, "the second replay runs one trace per thread");
    const TpChunk* rec = record(a, k, n0);
    const int status = rec->status;
    // A warp whose marked chunks all have the closed form writes them; if one of them has to be replayed
    // the others are replayed with it (from the same entry state: the same outputs) rather than making the
    // warp do both one after the other (random-walk depths, 4096 x 10^5: 2.82 -> 3.59 ms when it did).
    const unsigned active = __activemask();
    const bool replay = __any_sync(active, status == kRerun);
    if (!replay) {  // nothing to replay in this warp: its kFill chunks are written, here or by tp_fill_kernel
      const ksa_step_consts cc(a.pk);
      const int kind = g_tp_kind[(long long)k * a.N + n0];
      const unsigned fillable = __ballot_sync(active, tp_fillable(cc, status, kind));
      const bool group = ((fillable >> (threadIdx.x & 24)) & 0xffu) == 0xffu;  // my 8 adjacent traces
      if (status == kFill && !(group && tp_vector_layout(a))) tp_fill_chunk(a, cc, n0, k * a.L, min(a.L, a.T - k * a.L), kind, rec->entry1);
      return;
    }
    if (status != kRerun && status != kFill) return;
  }
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const int tb = (k == 0 || RERUN) ? t0 : t0 - a.warm;  // first tick replayed
  const int te = t0 + len;
  const ksa_step_consts c(a.pk);
  const int one = a.one;
  if (!RERUN && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_tp_count = 0;  // pass 2a's list starts empty
  ksa_trace_regs s[V];
#pragma unroll
  for (int v = 0; v < V; ++v) {
    s[v].r = a.st_r[n0 + v];
    if constexpr (RERUN) {
      const TpState e = record(a, k, n0 + v)->entry1;
      s[v] = ksa_trace_regs{e.r, e.su, e.sd};
    } else if (k == 0) {
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
  // the failure mask rides the same ring
  const unsigned char* fp = FAULTS ? a.fail + n0 : nullptr;
  unsigned fbuf[FAULTS ? D : 1][V == 8 ? 2 : 1];
  if constexpr (FAULTS) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      if (tb + j < te) tp_load_fail<V>(fp + (long long)(tb + j) * a.ld, fbuf[j]);
  }
  unsigned kacc[V];
#pragma unroll
  for (int v = 0; v < V; ++v) kacc[v] = 0xffffffffu;
  auto tick = [&](const int tt, const int j, const bool prefetch, const bool store) {
    long long msgs[V];
    unsigned fw[V == 8 ? 2 : 1];
    if constexpr (FAULTS) {
#pragma unroll
      for (int w = 0; w < (V == 8 ? 2 : 1); ++w) fw[w] = fbuf[j][w];
    }
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
      if constexpr (FAULTS) tp_load_fail<V>(fp + (long long)(tt + D) * a.ld, fbuf[j]);
    }
    int rr[V];
    unsigned ev[V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      if constexpr (FAULTS) {
        const unsigned f = V == 1 ? fw[0] : (fw[v / 4] >> (8 * (v % 4))) & 0xffu;  // a u8 load is zero-extended
        ev[v] = ksa_step_faults_fma2<false, false>(c, msgs[v], f, tt + 1, one, s[v]);
      } else {
        ev[v] = ksa_step_fma_byte(c, msgs[v], tt + 1, one, s[v]);
      }
      rr[v] = s[v].r;
      // the chunk's kind, gathered on the way (see below): x 21 brings the three lines of a direction (bits 0 / 2 /
      // 4 up, 1 / 3 / 5 down; at most one of each per tick, so nothing carries) onto bits 4 and 5
      if constexpr (!FAULTS && !RERUN)
        if (store) kacc[v] &= ev[v] * 21u;
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
  if constexpr (!RERUN) {
#pragma unroll
    for (int v = 0; v < V; ++v) record(a, k, n0 + v)->entry1 = TpState{s[v].r, s[v].su, s[v].sd};
  }
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
    if constexpr (!RERUN) record(a, k, n0 + v)->status = 0;
  }
  // Did EVERY tick of the chunk meet the scale-up threshold (kind 1), or the scale-down threshold (2)? A tick
  // that meets a threshold logs exactly one of that direction's three lines, whatever the state, so the
  // event bytes of this replay from a guess say it; 0 otherwise, and for policies whose thresholds overlap
  // or with a failure mask (a failed call takes no stamp). Such a chunk is the saturated queue (replicas at
  // max, ScaleUp called every cool-down period) or the idle one (at min, ScaleDown every period): the steady
  // states of a real fleet. Its calls are exactly one cool-down apart, their phase set by the first one, so
  // a chunk started from a guess deep inside such a stretch keeps the wrong phase for ever and no warm-up
  // helps; but what the chunk does to a state has a closed form (pure_exit), which is how pass 3a gets
  // through it. (A pass of its own over the event bytes took 0.29 ms at 4096 x 10^5.)
  if constexpr (!RERUN) {
    const bool ordered = a.pk.up_msgs > a.pk.down_msgs;
#pragma unroll
    for (int v = 0; v < V; ++v) {
      unsigned char kind = 0;
      if (!FAULTS && ordered) kind = (kacc[v] & 16u) ? 1 : ((kacc[v] & 32u) ? 2 : 0);
      g_tp_kind[(long long)k * a.N + n0 + v] = kind;
    }
  }
}

// Replay chunk (t0, len) from state s, rewriting the outputs, in groups of 8 ticks with the depths
// of the next group already in flight (a plain loop waits a memory latency per tick). Stops after a
// checkpoint tick j >= first_check whose pass-1 state has the same future; returns that j, or -1
// when the chunk ends first.
template <int A, bool FAULTS>
__device__ int tp_replay_until_match(const TpArgs& a, const ksa_step_consts& c, long long n, int t0, int len,
                                     const TpChunk* rec, int first_check, ksa_trace_regs& s) {
  constexpr int G = 8;
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + (long long)t0 * row + n;
  int* rp = a.replicas + (long long)t0 * a.ld + n;
  unsigned char* ep = a.events + (long long)t0 * a.ld + n;
  const unsigned char* fp = FAULTS ? a.fail + (long long)t0 * a.ld + n : nullptr;
  long long buf[G];
  unsigned fbuf[FAULTS ? G : 1];
#pragma unroll
  for (int j = 0; j < G; ++j) {
    buf[j] = j < len ? load_sum(dp + (long long)j * row, a.ld, A) : 0;
    if constexpr (FAULTS) fbuf[j] = j < len ? (unsigned)__ldg(fp + (long long)j * a.ld) : 0u;
  }
  for (int t = 0; t < len; t += G) {
#pragma unroll
    for (int j = 0; j < G; ++j) {
      const int tt = t + j;
      if (tt < len) {
        const long long msgs = buf[j];
        const unsigned f = FAULTS ? fbuf[j] : 0u;
        if (tt + G < len) {
          buf[j] = load_sum(dp + (long long)(tt + G) * row, a.ld, A);
          if constexpr (FAULTS) fbuf[j] = (unsigned)__ldg(fp + (long long)(tt + G) * a.ld);
        }
        const unsigned ev = FAULTS ? ksa_step_faults(c, msgs, f, t0 + tt + 1, s) : ksa_step(c, msgs, t0 + tt + 1, s);
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

// ---- before pass 1, for trace counts at which a second replay costs more than the plain kernel -----------
// Do guesses meet the truth on these depths at all? Where every tick meets the same threshold for long
// (a saturated or an idle queue) the scale calls come exactly one cool-down apart, and a run started
// from a guess keeps whatever phase its starting tick gave it: such chunks are replayed a second time
// (pass 3a), and from about 16384 traces on two chunked replays cost more than one plain one (random-walk
// depths, profiles/r02_small_n_modes.txt). The probe tells in 25 us: for a sample of traces and every
// chunk start, warm-ups of three lengths (their differences 61 and 97 ticks: no cool-down of a few
// ticks divides them) either end in states with the same future, as runs that have met the truth do,
// or not. Block = 32 (trace, chunk start) pairs x 3 warm-ups.
template <int A, bool FAULTS>
__global__ void __launch_bounds__(96) tp_probe_kernel(const __grid_constant__ TpArgs a) {
  __shared__ TpState ended[3][32];
  if (blockIdx.x == 0 && threadIdx.x == 0) g_tp_probe_bad = 0;
  const int variant = threadIdx.x / 32, lane = threadIdx.x % 32;
  const long long pair = (long long)blockIdx.x * 32 + lane;
  const long long pairs = (long long)a.probe_traces * (a.K - 1);
  const int k = (int)(pair / a.probe_traces) + 1;
  // the sample: groups of 32 adjacent traces (a warp's loads are one 256-byte run), the groups spread over the call's traces
  const long long sample = pair % a.probe_traces;
  const long long n = (sample / 32) * a.probe_stride + sample % 32;
  const bool live = pair < pairs && n < a.N;
  const int t0 = k * a.L;
  if (live) {
    const ksa_step_consts c(a.pk);
    const int w = min(variant == 0 ? 128 : (variant == 1 ? 189 : 286), t0);
    ksa_trace_regs s{a.st_r[n], 0, 0};  // pass 1's guess
    const int first = t0 - w;
    constexpr int G = 16;
    const long long row = (long long)A * a.ld;
    const long long* dp = a.depth + (long long)first * row + n;
    const unsigned char* fp = FAULTS ? a.fail + (long long)first * a.ld + n : nullptr;
    long long buf[G];
    unsigned fbuf[FAULTS ? G : 1];
#pragma unroll
    for (int j = 0; j < G; ++j) {
      buf[j] = j < w ? load_sum(dp + (long long)j * row, a.ld, A) : 0;
      if constexpr (FAULTS) fbuf[j] = j < w ? (unsigned)__ldg(fp + (long long)j * a.ld) : 0u;
    }
    for (int t = 0; t < w; t += G) {
#pragma unroll
      for (int j = 0; j < G; ++j) {
        const int tt = t + j;
        if (tt < w) {
          const long long msgs = buf[j];
          const unsigned f = FAULTS ? fbuf[j] : 0u;
          if (tt + G < w) {
            buf[j] = load_sum(dp + (long long)(tt + G) * row, a.ld, A);
            if constexpr (FAULTS) fbuf[j] = (unsigned)__ldg(fp + (long long)(tt + G) * a.ld);
          }
          if constexpr (FAULTS) ksa_step_faults(c, msgs, f, first + tt + 1, s);
          else ksa_step(c, msgs, first + tt + 1, s);
        }
      }
    }
    ended[variant][lane] = TpState{s.r, s.su, s.sd};
  }
  __syncthreads();
  if (live && variant == 0 &&
      !(same_future(ended[0][lane], ended[1][lane], t0) && same_future(ended[0][lane], ended[2][lane], t0)))
    atomicAdd(&g_tp_probe_bad, 1);
}

// ---- pass 2a -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(BLOCK) tp_check_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  const int k = blockIdx.y + 1;
  bool need = false;
  if (g_tp_probe_bad > a.probe_limit) return;
  if (n < a.N) {
    const TpState entry = record(a, k - 1, n)->exit1;
    TpChunk* const rec = record(a, k, n);
    need = !same_future(entry, rec->entry1, k * a.L);
    g_tp_delta[(long long)k * a.N + n] = 0;
    // only the replica count is off, and the chunk can be shifted by an offset: not for pass 2b. Pass 3
    // does that, with the true offset (which it alone knows: the previous chunk may itself get shifted)
    if (need && same_gates(entry, rec->entry1, k * a.L) && shift_safe(*rec)) {
      rec->status = kLeftToShift;
      g_tp_walk[n] = 1;
      need = false;
    }
    if (need) rec->status = kListed;  // pass 2b replaces it when it gets the chunk done
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
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_fix_kernel(const __grid_constant__ TpArgs a) {
  const int idx = blockIdx.x * BLOCK + threadIdx.x;
  // This pass is for the few chunks whose guess met the truth a little late. When many are listed the
  // guesses fail for a reason (stretches in which every tick meets a threshold: the phase of the calls):
  // everything is left to pass 3a and the second replay.
  if (idx >= g_tp_count || g_tp_probe_bad > a.probe_limit) return;
  const int item = g_tp_list[idx];
  const int k = (int)(item / a.N);
  const long long n = item % a.N;
  if (g_tp_count > a.fix_limit) {
    g_tp_walk[n] = 1;
    return;
  }
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const ksa_step_consts c(a.pk);
  TpChunk* const rec = record(a, k, n);
  const TpState entry = record(a, k - 1, n)->exit1;
  g_tp_used[item] = entry;
  ksa_trace_regs s{entry.r, entry.su, entry.sd};
  const int j = tp_replay_until_match<A, FAULTS>(a, c, n, t0, len, rec, 0, s);
  rec->entry1 = TpState{s.r, s.su, s.sd};  // (the slot is free: pass 2a has read it)
  rec->status = j >= 0 ? j + 1 : -1;
  if (j < 0) g_tp_walk[n] = 1;  // the chunk after it was judged by an exit state that is not the true one
}

// The state after a chunk of that kind: the first call at the first tick the gate is open, then one
// every cool-down period, each a full step towards the limit (scale.go:62-69, 89-97). Chunk = local
// ticks t0 + 1 .. t0 + len. False: not worked out (steps that are negative or could wrap).
__device__ __forceinline__ bool pure_exit(const ksa_step_consts& c, int kind, int t0, int len, TpState& g) {
  const bool up = kind == 1;
  const int step = up ? c.up32 : c.down32;
  if (!pure_steps_ok(c, kind)) return false;
  const int cool = up ? c.cu : c.cd;
  const long long first = max(t0 + 1, up ? g.su : g.sd);  // blocked at local tick i <=> i < s
  if (first > t0 + len) return true;                      // the gate stays shut for the whole chunk
  const long long period = max(cool, 1);
  const long long calls = (t0 + len - first) / period + 1;
  const long long last = first + (calls - 1) * period;
  if (up) {
    g.su = (int)(last + cool);
    if (g.r < c.max32) g.r = (int)min((long long)g.r + calls * step, (long long)c.max32);
  } else {
    g.sd = (int)(last + cool);
    if (g.r > c.min32) g.r = (int)max((long long)g.r - calls * step, (long long)c.min32);
  }
  return true;
}

// ---- pass 3 --------------------------------------------------------------------------------------
// A stamp value s was taken by a scale call after local tick t  <=>  s - c > t  (s = call tick + c).
__device__ __forceinline__ int later_stamp(int keep, int cand, int c, int after) {
  return (cand - c > after) ? cand : keep;
}

// The walk over a trace's chunks with the true state, for traces in which not every chunk simply stands.
//  - A chunk whose outputs are pass 1's or the second replay's (status 0, kLeftToShift, kListed, kRerun)
//    was replayed from entry1 and ended with exit1: it stands if entry1 has the true future, or, when only
//    the replica count is off and no scale call happened in it, after a shift (tp_shift_kernel).
//  - A chunk pass 2b patched (status j + 1, -1) was replayed from g_tp_used: it stands if that has the
//    true future.
//  Otherwise:
//  RESOLVE (pass 3a, before the second replay): the chunk is marked (kRerun) to be replayed a second time
//    from the true state, which goes into entry1. What it will end with is not known yet, and the next
//    chunk needs it: worked out for a chunk in which every tick met the same threshold (pure_exit), assumed
//    to be what it ended with the first time otherwise (most chunks forget the state they were entered
//    with). Wrong assumptions cost time, not correctness:
//  FINAL (pass 3b, after the second replay): the same walk over what is there now. A chunk that still does
//    not fit is replayed here, up to kChainReplays per trace; a trace that needs more is flagged for the
//    one-trace-per-thread kernel that follows and its state is left alone.
// Returns false if the trace is given up.
template <int A, bool FAULTS, bool RESOLVE>
__device__ bool tp_walk(const TpArgs& a, const ksa_step_consts& c, long long n, TpState& cur) {
  int replays = 0;
  // One thread walks a trace's chunks in order, and a record costs a memory latency: the next chunk's
  // (status, the two states, its kind) is on its way while this one is judged (4096 traces x 96 chunks:
  // a tenth of a millisecond per walk without that). Nothing below writes a later chunk's record.
  const TpChunk* ahead = record(a, 1, n);
  int status_next = ahead->status;
  TpState entry_next = ahead->entry1, exit_next = ahead->exit1;
  int kind_next = g_tp_kind[a.N + n];
  for (int k = 1; k < a.K; ++k) {
    const int t0 = k * a.L;
    const int len = min(a.L, a.T - t0);
    const long long idx = (long long)k * a.N + n;
    TpChunk* const rec = record(a, k, n);
    const int status = status_next;
    const TpState entry1 = entry_next, exit1 = exit_next;
    const int kind = kind_next;
    if (k + 1 < a.K) {
      ahead = record(a, k + 1, n);
      status_next = ahead->status;
      entry_next = ahead->entry1;
      exit_next = ahead->exit1;
      kind_next = g_tp_kind[idx + a.N];
    }
    const bool patched = status < 0 || (status >= 1 && status <= kChecks);
    if (!patched) {
      const long long delta = (long long)cur.r - entry1.r;
      if (same_gates(entry1, cur, t0) && (delta == 0 || (exit1.su == entry1.su && exit1.sd == entry1.sd))) {  // (shift_safe)
        if (!RESOLVE && delta != 0) g_tp_delta[idx] = (int)delta;
        cur.r = (int)(exit1.r + delta);
        cur.su = later_stamp(cur.su, exit1.su, c.cu, t0);
        cur.sd = later_stamp(cur.sd, exit1.sd, c.cd, t0);
        continue;
      }
    } else if (same_future(g_tp_used[idx], cur, t0)) {
      // pass 2b started from (the equivalent of) the true state: its work stands
      if (status < 0) {  // 2b rewrote the whole chunk; its end state is in the entry1 slot
        cur.r = entry1.r;
        cur.su = later_stamp(cur.su, entry1.su, c.cu, t0);
        cur.sd = later_stamp(cur.sd, entry1.sd, c.cd, t0);
      } else {
        const int tc = t0 + check_tick(status - 1);  // pass 1's trajectory is the true one after tc
        cur.su = later_stamp(cur.su, entry1.su, c.cu, t0);
        cur.sd = later_stamp(cur.sd, entry1.sd, c.cd, t0);
        cur.r = exit1.r;
        cur.su = later_stamp(cur.su, exit1.su, c.cu, tc);
        cur.sd = later_stamp(cur.sd, exit1.sd, c.cd, tc);
      }
      continue;
    }
    if (RESOLVE) {
      // what the chunk ended with the first time (pass 2b's end state is in the entry1 slot)
      const TpState ended = status < 0 ? entry1 : exit1;
      rec->entry1 = cur;
      rec->status = kRerun;
      TpState worked = cur;
      if (kind != 0 && pure_exit(c, kind, t0, len, worked)) {
        rec->status = kFill;  // the outputs have a closed form too: written, not replayed
        rec->exit1 = worked;
        cur = worked;
      } else {
        cur.r = ended.r;
        cur.su = later_stamp(cur.su, ended.su, c.cu, t0);
        cur.sd = later_stamp(cur.sd, ended.sd, c.cd, t0);
      }
    } else {
      if (++replays > kChainReplays) return false;  // left to the plain kernel that follows (g_tp_unresolved)
      ksa_trace_regs s{cur.r, cur.su, cur.sd};
      tp_replay_until_match<A, FAULTS>(a, c, n, t0, len, rec, kChecks, s);  // (the whole chunk: no checkpoint applies)
      cur = TpState{s.r, s.su, s.sd};
      g_tp_delta[idx] = 0;
      rec->status = kRedone;
    }
  }
  return true;
}

// ---- pass 3a: thread (trace) -----------------------------------------------------------------------
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_resolve_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N || g_tp_probe_bad > a.probe_limit || g_tp_walk[n] == 0) return;
  const ksa_step_consts c(a.pk);
  TpState cur = record(a, 0, n)->exit1;  // chunk 0 ran from the true state
  tp_walk<A, FAULTS, true>(a, c, n, cur);
}

// ---- pass 3b: thread (trace) -----------------------------------------------------------------------
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_chain_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  if (g_tp_probe_bad > a.probe_limit) {  // the probe called the chunked passes off: every trace is the plain kernel's
    g_tp_unresolved[n] = 1;
    return;
  }
  const ksa_step_consts c(a.pk);
  TpState cur = record(a, 0, n)->exit1;  // chunk 0 ran from the true state
  if (g_tp_walk[n] == 0) {
    // Every chunk stands as pass 1 wrote it, or as pass 2b patched it up to a checkpoint from the previous
    // chunk's exit (the common case): all entry states were the true ones, the replica count is the last
    // chunk's, and each stamp is the latest one taken where the trajectory was the true one. Valid
    // candidates are later than anything before them, so "latest" is a maximum and the loop has no
    // dependence from one chunk's record to the next one's loads.
    int su = cur.su, sd = cur.sd;
    for (int k = 1; k < a.K; ++k) {
      const TpChunk* rec = record(a, k, n);
      const int status = rec->status;
      const int t0 = k * a.L;
      const int tc = status == 0 ? t0 : t0 + check_tick(status - 1);
      if (status > 0) {  // pass 2b's own replay, from the chunk's first tick
        su = max(su, later_stamp(INT_MIN, rec->entry1.su, c.cu, t0));
        sd = max(sd, later_stamp(INT_MIN, rec->entry1.sd, c.cd, t0));
      }
      su = max(su, later_stamp(INT_MIN, rec->exit1.su, c.cu, tc));
      sd = max(sd, later_stamp(INT_MIN, rec->exit1.sd, c.cd, tc));
    }
    cur.r = record(a, a.K - 1, n)->exit1.r;
    cur.su = su;
    cur.sd = sd;
  } else {
    g_tp_walk[n] = 0;  // (for the next call)
    if (!tp_walk<A, FAULTS, false>(a, c, n, cur)) {
      g_tp_unresolved[n] = 1;
      return;
    }
  }
  g_tp_unresolved[n] = 0;
  a.st_r[n] = cur.r;
  a.st_ju[n] = ksa_global_stamp(cur.su, c.cu, a.tick0, a.st_ju[n]);
  a.st_jd[n] = ksa_global_stamp(cur.sd, c.cd, a.tick0, a.st_jd[n]);
}

// ---- after pass 3: the chunks it decided to shift --------------------------------------------------
// A chunk is shifted only if no scale call happened in it (shift_safe), so its replica count is ONE number
// from its first tick to its last: what it ended with, plus the offset. Where 8 adjacent traces all have such
// a chunk it is written, not read and added to: half the bytes, and 256-bit stores (the trace that goes
// quiet is a fleet that goes quiet: 131072 x 10^4 quiet traces, 4.30 ms with every count read and
// rewritten, the worst case against the plain kernels' 3.28). Other shifted chunks: tp_shift_kernel.
__device__ __forceinline__ bool tp_constant_chunk(const TpChunk& rec, int unresolved) {
  const int status = rec.status;
  const bool patched = status < 0 || (status >= 1 && status <= kChecks) || status == kRedone;  // (entry1 is not its entry state)
  return unresolved == 0 && !patched && shift_safe(rec);
}
// Thread (8 adjacent traces, chunk >= 1, part of the chunk).
__global__ void __launch_bounds__(BLOCK) tp_shift_fill_kernel(const __grid_constant__ TpArgs a) {
  const long long n0 = ((long long)blockIdx.x * BLOCK + threadIdx.x) * 8;
  if (n0 >= a.N || g_tp_probe_bad > a.probe_limit || !tp_vector_layout(a)) return;  // (N is a multiple of 8)
  const int k = blockIdx.y + 1;
  int r[8], delta[8];
  bool any = false, all = true;
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    delta[v] = g_tp_delta[(long long)k * a.N + n0 + v];
    any = any || delta[v] != 0;
  }
  if (!any) return;  // (the usual call stops here)
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    const TpChunk* rec = record(a, k, n0 + v);
    all = all && tp_constant_chunk(*rec, g_tp_unresolved[n0 + v]);
    r[v] = (int)((unsigned)rec->exit1.r + (unsigned)delta[v]);
  }
  if (!all) return;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const int part = (len + kFillParts - 1) / kFillParts;
  const int ta = min((int)blockIdx.z * part, len), tb = min(ta + part, len);
  int* rp = a.replicas + (long long)(t0 + ta) * a.ld + n0;
  for (int t = ta; t < tb; ++t) {
    asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(rp), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
    rp += a.ld;
  }
}
// Thread (trace, chunk >= 1): add the chunk's offset to its replica counts (the events are the same), unless
// tp_shift_fill_kernel had the chunk's group of 8 traces.
__global__ void __launch_bounds__(BLOCK) tp_shift_kernel(const __grid_constant__ TpArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const int k = blockIdx.y + 1;
  if (g_tp_probe_bad > a.probe_limit) return;
  const int delta = g_tp_delta[(long long)k * a.N + n];
  const int unresolved = g_tp_unresolved[n];
  const unsigned constant = __ballot_sync(__activemask(), tp_constant_chunk(*record(a, k, n), unresolved));
  if (delta == 0 || unresolved != 0) return;  // (a trace pass 3 gave up is rewritten by the plain kernel)
  if (tp_vector_layout(a) && ((constant >> (threadIdx.x & 24)) & 0xffu) == 0xffu) return;  // written already
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  int* rp = a.replicas + (long long)t0 * a.ld + n;
  constexpr int G = 8;  // loads of a group first: one memory latency per 8 ticks
  for (int t = 0; t < len; t += G) {
    int v[G];
#pragma unroll
    for (int j = 0; j < G; ++j) v[j] = t + j < len ? rp[(long long)(t + j) * a.ld] : 0;
#pragma unroll
    for (int j = 0; j < G; ++j)
      if (t + j < len) rp[(long long)(t + j) * a.ld] = (int)((unsigned)v[j] + (unsigned)delta);
  }
}

// ==== summary mode ================================================================================
//
// ksa_replay_summary / ksa_replay_summary_faults for few traces. A chunk of the time axis is
// replayed exactly like a call of its own: chunk k of trace n has a carry state (replicas and the
// two stamps as absolute ticks, the caller's own format) and a ksa_summary record in the scratch,
// and pass 1 IS the default summary kernel (ksa_replay.cu, one trace per thread) over a grid of
// (trace blocks, chunks). A record's counters, sum and min / max combine by addition and min / max;
// the fingerprint of a chunk started from 0 is an affine map of the fingerprint before it
// (lo' = lo * M^len + h), so the chunks' fingerprints chain by multiply-adds.
//
//  pass 0  thread (trace, chunk): chunk 0 takes the caller's state; chunk k > 0 replays the `warm`
//          ticks before its first one from a guess (the caller's state) and keeps what it ends
//          with as its entry state. Clears the chunk's summary record.
//  pass 1  the summary kernel: entry state -> exit state, summary record.
//  pass 2a thread (trace, chunk >= 1): does the previous chunk's exit state have the same future as
//          this chunk's entry state? If not, the chunk goes on the list.
//  pass 2b thread (list item): replay the whole chunk again from the previous chunk's exit state
//          (plain loop), overwriting its summary record; remember the state it ends with and whether
//          that has the same future as pass 1's exit state (which the next chunk was judged by).
//  pass 3  warp (trace): if every chunk is fine, all records were computed from the true states by
//          induction: the lanes combine them in order (each lane a run of chunks, then a tree over the
//          lanes). Otherwise lane 0 walks the chunks with the true state: a record replayed from the
//          right gates but a replica count that is off by an offset is shifted by it when the chunk
//          never came near a limit (tp_shift_safe: min, max, sum and the fingerprint's replica lane
//          move with the offset, everything else stands); up to two chunks that were replayed from a
//          wrong state are replayed again; a trace that needs more is flagged and left to the
//          one-trace-per-thread kernel that follows. The result is added to the caller's record and state.
//
// "Same future" at the first tick B + 1 of a chunk, with `rem` ticks of the call left: equal replica
// counts, and for each direction the same first tick at which the gate is open, both clamped into
// [1, rem + 1] (open from the chunk's first tick on is one case, shut for the rest of the call another).
// The caller's stamps are the true ones: per direction the latest stamp that a valid chunk took
// inside its own ticks (later than B), else what came before it.
//
// Scratch: the 64-byte slots of g_tp_scratch seen as arrays over (chunk, trace) — three carry
// states (pass 1's in/out, the entry copy, pass 2b's exit) and a status word — plus one ksa_summary per
// slot (g_tp_partial, 32 MB).

static __device__ __align__(256) ksa_summary g_tp_partial[kMaxChunks];

struct TpCarry {
  int* r;
  long long* ju;
  long long* jd;
};
struct TpSumArgs {
  ksa_policy_k pk;
  const long long* depth;
  const unsigned char* fail;
  long long ld, tick0, N;
  int T, A, K, L, warm, one;
  int piece;  // pass 2b cuts a chunk into pieces of this many ticks (a multiple of 8, at most 32 pieces)
  int fix_limit;  // ... unless more chunks than this are listed
  unsigned mL_lo, mL_hi, mE_lo, mE_hi;  // the fingerprint multipliers to the power L / the last chunk's length
  unsigned gL_lo, gE_lo;                // 1 + M + ... + M^(len-1) for the replica lane's multiplier, same two lengths
  int* st_r;           // the caller's state and summaries
  long long* st_ju;
  long long* st_jd;
  ksa_summary* summary;
  TpCarry io, en, fx;  // scratch: pass 1's state (in: entry, out: exit), entry copy, pass 2b's exit
  int* status;         // 0 pass 1's record stands; 1 pass 2b's record, exit as pass 1's; -1 pass 2b's, exit differs;
                       // 2 pass 1's record, entry off by a replica offset the chunk can be shifted by (pass 3 does);
                       // 3 replayed a second time from the entry state pass 3a worked out (in `en`); 4 pass 1's
                       // record of a chunk in which every tick met the same threshold, left to pass 3a; 5 pass 1's
                       // record of a chunk listed for pass 2b (which replaces the 5 unless it gives the chunk up);
                       // 6 record and exit state worked out by pass 3a from the entry state in `en` (tp_pure_record;
                       // 7 / 8 on the way there)
  ksa_summary* part;
};

struct TpG {  // a carry state in the caller's format
  int r;
  long long ju, jd;
};
__device__ __forceinline__ TpG tp_get(const TpCarry& c, long long i) { return TpG{c.r[i], c.ju[i], c.jd[i]}; }
__device__ __forceinline__ void tp_put(const TpCarry& c, long long i, const TpG& g) {
  c.r[i] = g.r;
  c.ju[i] = g.ju;
  c.jd[i] = g.jd;
}
// First tick after B (1-based) at which a gate with stamp `last` is open, clamped into [1, rem + 1].
__device__ __forceinline__ int tp_gate(long long cool, int c, long long B, long long last, int rem) {
  return min(max(ksa_local_stamp(cool, c, B, last), 1), rem + 1);
}
__device__ __forceinline__ bool tp_same_gates(const ksa_policy_k& pk, const ksa_step_consts& c, const TpG& a,
                                              const TpG& b, long long B, int rem) {
  return tp_gate(pk.cu, c.cu, B, a.ju, rem) == tp_gate(pk.cu, c.cu, B, b.ju, rem) &&
         tp_gate(pk.cd, c.cd, B, a.jd, rem) == tp_gate(pk.cd, c.cd, B, b.jd, rem);
}
__device__ __forceinline__ bool tp_same_future(const ksa_policy_k& pk, const ksa_step_consts& c, const TpG& a,
                                               const TpG& b, long long B, int rem) {
  return a.r == b.r && tp_same_gates(pk, c, a, b, B, rem);
}

// A chunk replayed from the right gates but a replica count that was off by `delta` (true = replayed +
// delta) need not be replayed again if it never came near a limit: which calls happen does not depend
// on the replica count (a call that finds the limit reached stamps like any other, main.go:62,76;
// scale.go:62-65), so both runs make the same calls, and as long as every one of them is applied in
// full in both — no clamp (scale.go:67-69, 95-97), no call at the limit — the true trajectory is the
// replayed one plus delta at every tick. Sufficient, from the chunk's record `q` (replayed from
// `entry_r`): no scale-up call at all, or every count of the chunk below max in both runs (a clamped or
// refused step leaves the count at or above max); likewise for scale-down. The chunk that goes quiet
// after some activity — no call for thousands of ticks — is the case this is for: its state never
// forgets, so no amount of warm-up makes a guess meet it.
__device__ __forceinline__ bool tp_shift_safe(const ksa_step_consts& c, const ksa_summary& q, int entry_r,
                                              long long delta) {
  const long long hi = max((long long)q.max_replicas, (long long)entry_r);
  const long long lo = min((long long)q.min_replicas, (long long)entry_r);
  if (hi + delta > INT_MAX || lo + delta < INT_MIN) return false;
  // With overlapping thresholds a tick can step up AND down: the count after the tick does not show
  // whether the step up was clamped in between (found by tools/fuzz_timepar.py). Only chunks without any
  // call are shifted then.
  if (!(c.up_msgs > c.down_msgs) && (q.up_applied | q.up_noop | q.down_applied | q.down_noop) != 0u) return false;
  if ((q.up_applied | q.up_noop) != 0u &&
      !(c.up32 >= 0 && c.up32 <= (1 << 30) && hi < c.max32 && hi + delta < c.max32 && step_cannot_wrap(c, true)))
    return false;
  if ((q.down_applied | q.down_noop) != 0u &&
      !(c.down32 >= 0 && c.down32 <= (1 << 30) && lo > c.min32 && lo + delta > c.min32 && step_cannot_wrap(c, false)))
    return false;
  return true;
}
// The record and exit state of such a chunk (`len` ticks; g_lo = 1 + M + ... + M^(len-1) for the
// fingerprint's replica lane: every tick's term moves by delta).
__device__ __forceinline__ void tp_shift(ksa_summary& q, TpG& exit, long long delta, int len, unsigned g_lo) {
  q.final_replicas = (int)(q.final_replicas + delta);
  q.min_replicas = (int)(q.min_replicas + delta);
  q.max_replicas = (int)(q.max_replicas + delta);
  q.sum_replicas = (long long)((unsigned long long)q.sum_replicas + (unsigned long long)delta * (unsigned long long)len);
  const unsigned lo = (unsigned)q.hash + (unsigned)delta * g_lo;
  q.hash = (q.hash & 0xffffffff00000000ull) | lo;
  exit.r = (int)(exit.r + delta);
}

__device__ __forceinline__ void tp_summary_clear(ksa_summary& q) {
  q.final_replicas = 0;
  q.min_replicas = INT_MAX;
  q.max_replicas = INT_MIN;
  q.up_applied = q.down_applied = q.up_noop = q.down_noop = q.up_cool = q.down_cool = 0u;
  q.poll_failed = q.up_failed = q.down_failed = 0u;
  q.sum_replicas = 0;
  q.hash = 0ull;
}
// One tick into a record, the plain way (the definition of the fields, include/ksa_replay.h).
__device__ __forceinline__ void tp_summary_add(ksa_summary& q, int r, unsigned ev) {
  q.final_replicas = r;
  q.min_replicas = min(q.min_replicas, r);
  q.max_replicas = max(q.max_replicas, r);
  q.up_applied += (ev & KSA_EV_UP_APPLIED) ? 1u : 0u;
  q.down_applied += (ev & KSA_EV_DOWN_APPLIED) ? 1u : 0u;
  q.up_noop += (ev & KSA_EV_UP_NOOP) ? 1u : 0u;
  q.down_noop += (ev & KSA_EV_DOWN_NOOP) ? 1u : 0u;
  q.up_cool += (ev & KSA_EV_UP_COOL) ? 1u : 0u;
  q.down_cool += (ev & KSA_EV_DOWN_COOL) ? 1u : 0u;
  const unsigned fb = ev & (unsigned)KSA_EV_POLL_FAILED;  // both failure bits = the poll failed
  q.poll_failed += fb == (unsigned)KSA_EV_POLL_FAILED ? 1u : 0u;
  q.up_failed += fb == (unsigned)KSA_EV_UP_FAILED ? 1u : 0u;
  q.down_failed += fb == (unsigned)KSA_EV_DOWN_FAILED ? 1u : 0u;
  q.sum_replicas = (long long)((unsigned long long)q.sum_replicas + (unsigned long long)(long long)r);
  q.hash = ksa_hash_step(q.hash, r, ev);
}

// A chunk in which EVERY tick met the scale-up threshold (1) or the scale-down threshold (2), from its
// record (a tick that meets a threshold logs exactly one of that direction's three lines); 0 otherwise,
// and for policies whose thresholds overlap (both can be met in one tick) or with a failure mask (a
// failed call takes no stamp). That is a property of the depths, not of the state the chunk was replayed
// from. It is the saturated queue (replicas at max, ScaleUp called every cool-down period) and the idle
// one (at min, ScaleDown every period): the two steady states of a real fleet. The calls of such a
// stretch are exactly one cool-down apart, their phase set by its first call, so a chunk started from
// a guess deep inside one keeps the wrong phase for ever: no warm-up helps.
template <bool FAULTS>
__device__ __forceinline__ int tp_pure_kind(const ksa_policy_k& pk, const ksa_summary& q, int len) {
  if (FAULTS || !(pk.up_msgs > pk.down_msgs)) return 0;
  if ((long long)q.up_applied + q.up_noop + q.up_cool == len) return 1;
  if ((long long)q.down_applied + q.down_noop + q.down_cool == len) return 2;
  return 0;
}
// ... but what such a chunk does to a state is simple: the first call at the first tick the gate is
// open, then one every cool-down period, each a full step towards the limit. The state after the
// chunk's `len` ticks (first tick B + 1) from state g. False: not worked out (steps that are negative
// or could wrap).
__device__ __forceinline__ bool tp_pure_exit(const ksa_policy_k& pk, const ksa_step_consts& c, int kind, long long B,
                                             int len, TpG& g) {
  const bool up = kind == 1;
  const int step = up ? c.up32 : c.down32;
  if (step < 0 || step > (1 << 30) || !step_cannot_wrap(c, up)) return false;
  const int cool = up ? c.cu : c.cd;
  const int s = ksa_local_stamp(up ? pk.cu : pk.cd, cool, B, up ? g.ju : g.jd);  // blocked at local tick i <=> i < s
  const long long first = max(1, s);
  if (first > len) return true;  // the gate stays shut for the whole chunk
  const long long period = max(cool, 1);
  const long long calls = (len - first) / period + 1;
  const long long last = first + (calls - 1) * period;
  if (up) {
    g.ju = B + last;
    if (g.r < c.max32) g.r = (int)min((long long)g.r + calls * step, (long long)c.max32);  // scale.go:62-69
  } else {
    g.jd = B + last;
    if (g.r > c.min32) g.r = (int)max((long long)g.r - calls * step, (long long)c.min32);  // scale.go:89-97
  }
  return true;
}

// m^n and 1 + m + ... + m^(n-1) mod 2^32 (bits of n from the top: n -> 2n: G (1 + P), P P; -> 2n + 1: G m + 1, P m)
__device__ __forceinline__ void tp_pow_geo(unsigned m, unsigned n, unsigned& pw, unsigned& geo) {
  pw = 1u;
  geo = 0u;
  for (int b = 31 - __clz((int)(n | 1u)); b >= 0; --b) {
    geo *= 1u + pw;
    pw *= pw;
    if ((n >> b) & 1u) {
      geo = geo * m + 1u;
      pw *= m;
    }
  }
}
// ... and the whole RECORD of such a chunk has a closed form too, so that the saturated and the idle
// queue cost no second replay: ticks before the first call and between calls log the cool-down line,
// the first calls step the count towards the limit (at most 64 distinct counts are followed here, else
// false: the chunk is replayed), the rest find the limit reached. Sum, min / max and the counters are
// sums over those runs of equal counts; the fingerprint's replica lane is sum(count_t M^(len - t)), a
// geometric sum per run, and its event lane the cool-down code's geometric sum plus corrections at the
// call ticks, which form an arithmetic progression of exponents: a geometric sum with ratio M^period.
// `in` = the true state before the chunk's first tick (B + 1); fills q and `out`.
__device__ bool tp_pure_record(const ksa_policy_k& pk, const ksa_step_consts& c, int kind, long long B, int len,
                               const TpG& in, ksa_summary& q, TpG& out) {
  const bool up = kind == 1;
  const int step = up ? c.up32 : c.down32;
  if (step < 1 || step > (1 << 30) || !step_cannot_wrap(c, up)) return false;
  const int cool = up ? c.cu : c.cd;
  const int s = ksa_local_stamp(up ? pk.cu : pk.cd, cool, B, up ? in.ju : in.jd);  // blocked at local tick i <=> i < s
  const long long first = max(1, s), period = max(cool, 1);
  const long long calls = first > len ? 0 : (len - first) / period + 1;
  const long long limit = up ? c.max32 : c.min32;
  long long applied = 0;  // calls that still change the count
  if (up ? in.r < limit : in.r > limit) applied = ((up ? limit - in.r : (long long)in.r - limit) + step - 1) / step;
  applied = min(applied, calls);
  if (applied > 64) return false;
  const unsigned w_cool = up ? KSA_W_UP_COOL : KSA_W_DOWN_COOL, w_applied = up ? KSA_W_UP_APPLIED : KSA_W_DOWN_APPLIED,
                 w_noop = up ? KSA_W_UP_NOOP : KSA_W_DOWN_NOOP;
  tp_summary_clear(q);
  unsigned lo = 0u, hi = 0u, pw, geo;
  // a run of equal counts: local ticks [t0, t1]
  auto run = [&](long long value, long long t0, long long t1) {
    if (t1 < t0) return;
    const int r = (int)value;
    q.final_replicas = r;
    q.min_replicas = min(q.min_replicas, r);
    q.max_replicas = max(q.max_replicas, r);
    q.sum_replicas = (long long)((unsigned long long)q.sum_replicas + (unsigned long long)(long long)r * (unsigned long long)(t1 - t0 + 1));
    unsigned tail, unused;
    tp_pow_geo(KSA_HASH_MUL_LO, (unsigned)(len - t1), tail, unused);
    tp_pow_geo(KSA_HASH_MUL_LO, (unsigned)(t1 - t0 + 1), pw, geo);
    lo += (unsigned)r * tail * geo;
  };
  long long r = in.r;
  run(r, 1, calls == 0 ? len : first - 1);
  for (long long j = 1; j <= applied; ++j) {
    r = up ? min(r + step, limit) : max(r - step, limit);
    const long long t = first + (j - 1) * period;
    run(r, t, j == calls ? len : t + period - 1);
    tp_pow_geo(KSA_HASH_MUL_HI, (unsigned)(len - t), pw, geo);
    hi += (w_applied - w_cool) * pw;
  }
  if (applied < calls) {
    run(r, first + applied * period, len);
    // the calls that find the limit reached: ticks first + j period, j = applied .. calls - 1
    const long long last = first + (calls - 1) * period;
    unsigned ratio, tail;
    tp_pow_geo(KSA_HASH_MUL_HI, (unsigned)period, ratio, geo);
    tp_pow_geo(KSA_HASH_MUL_HI, (unsigned)(len - last), tail, geo);
    tp_pow_geo(ratio, (unsigned)(calls - applied), pw, geo);
    hi += (w_noop - w_cool) * tail * geo;
  }
  tp_pow_geo(KSA_HASH_MUL_HI, (unsigned)len, pw, geo);
  hi += w_cool * geo;
  q.hash = ((unsigned long long)hi << 32) | lo;
  (up ? q.up_applied : q.down_applied) = (unsigned)applied;
  (up ? q.up_noop : q.down_noop) = (unsigned)(calls - applied);
  (up ? q.up_cool : q.down_cool) = (unsigned)(len - calls);
  out = in;
  out.r = (int)r;
  if (calls > 0) (up ? out.ju : out.jd) = B + first + (calls - 1) * period;
  return true;
}

// Ticks [first, first + count) of the call for trace n, from carry state g (plain loop, the depths
// of the next 8 ticks in flight). With `q` the ticks are added to that record.
template <int A, bool FAULTS>
__device__ void tp_sum_replay(const TpSumArgs& a, const ksa_step_consts& c, long long n, int first, int count, TpG& g,
                              ksa_summary* q) {
  constexpr int G = 8;
  const long long B = a.tick0 + first;
  ksa_trace_regs s{g.r, ksa_local_stamp(a.pk.cu, c.cu, B, g.ju), ksa_local_stamp(a.pk.cd, c.cd, B, g.jd)};
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + (long long)first * row + n;
  const unsigned char* fp = FAULTS ? a.fail + (long long)first * a.ld + n : nullptr;
  long long buf[G];
  unsigned fbuf[FAULTS ? G : 1];
#pragma unroll
  for (int j = 0; j < G; ++j) {
    buf[j] = j < count ? load_sum(dp + (long long)j * row, a.ld, A) : 0;
    if constexpr (FAULTS) fbuf[j] = j < count ? (unsigned)__ldg(fp + (long long)j * a.ld) : 0u;
  }
  for (int t = 0; t < count; t += G) {
#pragma unroll
    for (int j = 0; j < G; ++j) {
      const int tt = t + j;
      if (tt < count) {
        const long long msgs = buf[j];
        const unsigned f = FAULTS ? fbuf[j] : 0u;
        if (tt + G < count) {
          buf[j] = load_sum(dp + (long long)(tt + G) * row, a.ld, A);
          if constexpr (FAULTS) fbuf[j] = (unsigned)__ldg(fp + (long long)(tt + G) * a.ld);
        }
        const unsigned ev = FAULTS ? ksa_step_faults(c, msgs, f, tt + 1, s) : ksa_step(c, msgs, tt + 1, s);
        if (q) tp_summary_add(*q, s.r, ev);
      }
    }
  }
  g.r = s.r;
  g.ju = ksa_global_stamp(s.su, c.cu, B, g.ju);
  g.jd = ksa_global_stamp(s.sd, c.cd, B, g.jd);
}

// `count` ticks (a multiple of 8) from tick `first` of the call for trace n, nothing recorded: the
// warm-up of a speculative start. The FMA-steered steps of the main kernels, the depths of the next
// 8 ticks in flight.
template <int A, bool FAULTS>
__device__ __forceinline__ void tp_sum_warm(const TpSumArgs& a, const ksa_step_consts& c, long long n, int first,
                                            int count, TpG& g) {
  constexpr int G = 8;
  const long long B = a.tick0 + first;
  ksa_trace_regs s{g.r, ksa_local_stamp(a.pk.cu, c.cu, B, g.ju), ksa_local_stamp(a.pk.cd, c.cd, B, g.jd)};
  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + (long long)first * row + n;
  const unsigned char* fp = FAULTS ? a.fail + (long long)first * a.ld + n : nullptr;
  const int one = a.one;
  long long buf[G][A];
  unsigned fbuf[FAULTS ? G : 1][1];
#pragma unroll
  for (int j = 0; j < G; ++j) {
#pragma unroll
    for (int p = 0; p < A; ++p) tp_load<1>(dp + (long long)j * row + p * a.ld, *(long long(*)[1]) & buf[j][p]);
    if constexpr (FAULTS) tp_load_fail<1>(fp + (long long)j * a.ld, fbuf[j]);
  }
#pragma unroll 1
  for (int t = 0; t < count; t += G) {
    const bool more = t + G < count;
#pragma unroll
    for (int j = 0; j < G; ++j) {
      unsigned long long acc = (unsigned long long)buf[j][0];  // sqs/sqs.go:56-64, wrapping
#pragma unroll
      for (int p = 1; p < A; ++p) acc += (unsigned long long)buf[j][p];
      const unsigned f = FAULTS ? fbuf[j][0] : 0u;
      if (more) {
#pragma unroll
        for (int p = 0; p < A; ++p)
          tp_load<1>(dp + (long long)(t + j + G) * row + p * a.ld, *(long long(*)[1]) & buf[j][p]);
        if constexpr (FAULTS) tp_load_fail<1>(fp + (long long)(t + j + G) * a.ld, fbuf[j]);
      }
      if constexpr (FAULTS) ksa_step_faults_fma2<false, false>(c, (long long)acc, f, t + j + 1, one, s);
      else ksa_step_fma_byte(c, (long long)acc, t + j + 1, one, s);
    }
  }
  g.r = s.r;
  g.ju = ksa_global_stamp(s.su, c.cu, B, g.ju);
  g.jd = ksa_global_stamp(s.sd, c.cd, B, g.jd);
}

// ---- pass 0 --------------------------------------------------------------------------------------
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_sum_warm_kernel(const __grid_constant__ TpSumArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_tp_count = 0;  // pass 2a's list starts empty
  if (n >= a.N) return;
  const int k = blockIdx.y;
  const ksa_step_consts c(a.pk);
  TpG g{a.st_r[n], a.st_ju[n], a.st_jd[n]};
  if (k > 0 && a.warm > 0) tp_sum_warm<A, FAULTS>(a, c, n, k * a.L - a.warm, a.warm, g);
  const long long idx = (long long)k * a.N + n;
  tp_put(a.io, idx, g);
  tp_put(a.en, idx, g);
  a.status[idx] = 0;
  ksa_summary q;
  tp_summary_clear(q);
  a.part[idx] = q;
}

// ---- pass 2a -------------------------------------------------------------------------------------
template <bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_sum_check_kernel(const __grid_constant__ TpSumArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  const int k = blockIdx.y + 1;
  bool need = false;
  if (n < a.N) {
    const ksa_step_consts c(a.pk);
    const long long idx = (long long)k * a.N + n;
    const TpG entry = tp_get(a.io, idx - a.N), guessed = tp_get(a.en, idx);
    const long long B = a.tick0 + (long long)k * a.L;
    need = !tp_same_future(a.pk, c, entry, guessed, B, a.T - k * a.L);
    // only the replica count is off, and by an offset the chunk can be shifted by: not for pass 2b. Pass 3
    // does that, with the true offset (which it alone knows: the previous chunk may itself be shifted)
    if (need && tp_same_gates(a.pk, c, entry, guessed, B, a.T - k * a.L) &&
        tp_shift_safe(c, a.part[idx], guessed.r, (long long)entry.r - guessed.r)) {
      a.status[idx] = 2;
      need = false;
    }
    // every tick of the chunk met the same threshold: replaying it in pieces would meet the same wrong
    // phases. Pass 3a works out its true entry state, and it is replayed a second time from that
    if (need && tp_pure_kind<FAULTS>(a.pk, a.part[idx], min(a.L, a.T - k * a.L)) != 0) {
      a.status[idx] = 4;
      need = false;
    }
    if (need) a.status[idx] = 5;  // listed; pass 2b replaces it when it gets the chunk done
  }
  const unsigned vote = __ballot_sync(0xffffffffu, need);  // one atomic per warp
  if (vote == 0u) return;
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == __ffs(vote) - 1) base = atomicAdd(&g_tp_count, __popc(vote));
  base = __shfl_sync(0xffffffffu, base, __ffs(vote) - 1);
  if (need) g_tp_list[base + __popc(vote & ((1u << lane) - 1u))] = (int)((long long)k * a.N + n);
}

// The stamp was taken inside ticks (B, B + len] (a stamp is the absolute tick it was taken at).
__device__ __forceinline__ bool tp_taken(long long stamp, long long B, int len) { return stamp > B && stamp <= B + len; }

// What a run of consecutive chunks adds up to. `m`, `h`: the run maps a fingerprint lane x to x * m + h.
struct TpRun {
  int any;  // the run has at least one chunk
  int r;    // replicas after its last tick
  int has_u, has_d;  // a stamp was taken inside the run
  long long ju, jd;
  int rmin, rmax;
  unsigned cnt[9];
  long long sum;
  unsigned m_lo, h_lo, m_hi, h_hi;
};
__device__ __forceinline__ void tp_run_init(TpRun& x) {
  x.any = x.r = x.has_u = x.has_d = 0;
  x.ju = x.jd = 0;
  x.rmin = INT_MAX;
  x.rmax = INT_MIN;
#pragma unroll
  for (int i = 0; i < 9; ++i) x.cnt[i] = 0u;
  x.sum = 0;
  x.m_lo = x.m_hi = 1u;
  x.h_lo = x.h_hi = 0u;
}
// x followed by y
__device__ __forceinline__ void tp_run_append(TpRun& x, const TpRun& y) {
  if (y.any) {
    x.any = 1;
    x.r = y.r;
  }
  if (y.has_u) {
    x.has_u = 1;
    x.ju = y.ju;
  }
  if (y.has_d) {
    x.has_d = 1;
    x.jd = y.jd;
  }
  x.rmin = min(x.rmin, y.rmin);
  x.rmax = max(x.rmax, y.rmax);
#pragma unroll
  for (int i = 0; i < 9; ++i) x.cnt[i] += y.cnt[i];
  x.sum = (long long)((unsigned long long)x.sum + (unsigned long long)y.sum);
  x.h_lo = x.h_lo * y.m_lo + y.h_lo;
  x.m_lo *= y.m_lo;
  x.h_hi = x.h_hi * y.m_hi + y.h_hi;
  x.m_hi *= y.m_hi;
}
// One chunk as a run: its record, the state it ended with, the tick before its first (B), its
// length, and whether its stamps count wherever they were taken (chunk 0 carries the caller's own).
__device__ __forceinline__ void tp_run_of_chunk(TpRun& x, const ksa_summary& q, const TpG& exit, long long B, int len,
                                                bool whole, unsigned m_lo, unsigned m_hi) {
  x.any = 1;
  x.r = exit.r;
  x.has_u = whole || tp_taken(exit.ju, B, len);
  x.has_d = whole || tp_taken(exit.jd, B, len);
  x.ju = exit.ju;
  x.jd = exit.jd;
  x.rmin = q.min_replicas;
  x.rmax = q.max_replicas;
  x.cnt[0] = q.up_applied;
  x.cnt[1] = q.down_applied;
  x.cnt[2] = q.up_noop;
  x.cnt[3] = q.down_noop;
  x.cnt[4] = q.up_cool;
  x.cnt[5] = q.down_cool;
  x.cnt[6] = q.poll_failed;
  x.cnt[7] = q.up_failed;
  x.cnt[8] = q.down_failed;
  x.sum = q.sum_replicas;
  x.m_lo = m_lo;
  x.h_lo = (unsigned)q.hash;
  x.m_hi = m_hi;
  x.h_hi = (unsigned)(q.hash >> 32);
}
__device__ __forceinline__ long long tp_shfl_ll(long long v, int src) {
  const int lo = __shfl_sync(0xffffffffu, (int)(unsigned)v, src);
  const int hi = __shfl_sync(0xffffffffu, (int)((unsigned long long)v >> 32), src);
  return (long long)(((unsigned long long)(unsigned)hi << 32) | (unsigned)lo);
}
__device__ __forceinline__ void tp_run_shfl(TpRun& y, const TpRun& x, int src) {
  y.any = __shfl_sync(0xffffffffu, x.any, src);
  y.r = __shfl_sync(0xffffffffu, x.r, src);
  y.has_u = __shfl_sync(0xffffffffu, x.has_u, src);
  y.has_d = __shfl_sync(0xffffffffu, x.has_d, src);
  y.ju = tp_shfl_ll(x.ju, src);
  y.jd = tp_shfl_ll(x.jd, src);
  y.rmin = __shfl_sync(0xffffffffu, x.rmin, src);
  y.rmax = __shfl_sync(0xffffffffu, x.rmax, src);
#pragma unroll
  for (int i = 0; i < 9; ++i) y.cnt[i] = __shfl_sync(0xffffffffu, x.cnt[i], src);
  y.sum = tp_shfl_ll(x.sum, src);
  y.m_lo = __shfl_sync(0xffffffffu, x.m_lo, src);
  y.h_lo = __shfl_sync(0xffffffffu, x.h_lo, src);
  y.m_hi = __shfl_sync(0xffffffffu, x.m_hi, src);
  y.h_hi = __shfl_sync(0xffffffffu, x.h_hi, src);
}

__device__ __forceinline__ TpG tp_shfl_g(const TpG& g, int src) {
  return TpG{__shfl_sync(0xffffffffu, g.r, src), tp_shfl_ll(g.ju, src), tp_shfl_ll(g.jd, src)};
}
__device__ __forceinline__ unsigned tp_pow_u32(unsigned base, unsigned e) {
  unsigned r = 1u;
  for (; e; e >>= 1, base *= base)
    if (e & 1u) r *= base;
  return r;
}
// ---- pass 2b -------------------------------------------------------------------------------------
// A listed chunk has to be replayed from its true entry state. One thread doing that alone needs the
// whole chunk's ticks one after the other (a millisecond for 1350 ticks: twice pass 1). A warp does
// it the way the call itself is done, one level down: the chunk is cut into up to 32 pieces of
// a.piece ticks, lane 0 replays the first from the true entry state, every other lane warms up from
// the guess before its piece and replays it; then the boundaries are checked in order, and a lane
// whose entry state did not have the true future replays its piece again from its left neighbour's
// exit state. Usually no boundary fails: warm-up + piece ticks of latency instead of the chunk's.
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_sum_fix_kernel(const __grid_constant__ TpSumArgs a) {
  const int item = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
  // This pass is for the few chunks whose guess met the truth a little late. When many are listed the
  // guesses fail for a reason pieces would fail for too (profiles/r02_small_n_modes.txt: random-walk
  // depths, 10.6 ms in here): everything is left to pass 3a and the second replay.
  if (item >= g_tp_count || g_tp_count > a.fix_limit) return;  // (whole warps)
  const int lane = threadIdx.x & 31;
  const long long idx = g_tp_list[item];
  const int k = (int)(idx / a.N);
  const long long n = idx % a.N;
  const int t0 = k * a.L;
  const int len = min(a.L, a.T - t0);
  const ksa_step_consts c(a.pk);
  const int pieces = (len + a.piece - 1) / a.piece;  // <= 32
  const bool mine = lane < pieces;
  const int first = t0 + lane * a.piece;             // my piece: ticks [first, first + count) of the call
  const int count = mine ? min(a.piece, t0 + len - first) : 0;
  const TpG entry0 = tp_get(a.io, idx - a.N);  // the previous chunk's exit state from pass 1: the chunk's entry
  TpG g = entry0, used = entry0;
  ksa_summary q;
  tp_summary_clear(q);
  if (mine) {
    if (lane > 0) {
      g = TpG{a.st_r[n], a.st_ju[n], a.st_jd[n]};
      const int w = min(a.warm, first);
      if (w > 0) tp_sum_warm<A, FAULTS>(a, c, n, first - w, w, g);
      used = g;
    }
    tp_sum_replay<A, FAULTS>(a, c, n, first, count, g, &q);
  }
  int redone = 0;
  for (int r = 0; r + 1 < pieces; ++r) {
    const TpG left = tp_shfl_g(g, r);  // exit state of piece r: true by induction
    bool redo = false;
    if (lane == r + 1) redo = !tp_same_future(a.pk, c, left, used, a.tick0 + first, a.T - first);
    if (__shfl_sync(0xffffffffu, (int)redo, r + 1)) {
      if (++redone > 2) return;  // the pieces' guesses keep failing: nothing written, the chunk is pass 3a's
      if (lane == r + 1) {
        g = used = left;
        tp_summary_clear(q);
        tp_sum_replay<A, FAULTS>(a, c, n, first, count, g, &q);
      }
    }
  }
  // the pieces in order -> the chunk
  TpRun run;
  tp_run_init(run);
  if (mine) {
    const long long B = a.tick0 + first;
    tp_run_of_chunk(run, q, g, B, count, false, tp_pow_u32(KSA_HASH_MUL_LO, (unsigned)count),
                    tp_pow_u32(KSA_HASH_MUL_HI, (unsigned)count));
  }
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    TpRun other;
    tp_run_shfl(other, run, min(lane + off, 31));
    if ((lane & (2 * off - 1)) == 0 && lane + off < 32) tp_run_append(run, other);
  }
  if (lane != 0) return;
  q.final_replicas = run.r;
  q.min_replicas = run.rmin;
  q.max_replicas = run.rmax;
  q.up_applied = run.cnt[0];
  q.down_applied = run.cnt[1];
  q.up_noop = run.cnt[2];
  q.down_noop = run.cnt[3];
  q.up_cool = run.cnt[4];
  q.down_cool = run.cnt[5];
  q.poll_failed = run.cnt[6];
  q.up_failed = run.cnt[7];
  q.down_failed = run.cnt[8];
  q.sum_replicas = run.sum;
  q.hash = ((unsigned long long)run.h_hi << 32) | run.h_lo;
  a.part[idx] = q;
  const TpG exit{run.r, run.has_u ? run.ju : entry0.ju, run.has_d ? run.jd : entry0.jd};
  tp_put(a.fx, idx, exit);
  tp_put(a.en, idx, entry0);  // what the record was replayed from (pass 3 compares it with the true state)
  const bool last = k == a.K - 1;  // (nothing is judged by the last chunk's exit)
  a.status[idx] =
      last || tp_same_future(a.pk, c, exit, tp_get(a.io, idx), a.tick0 + t0 + len, a.T - t0 - len) ? 1 : -1;
}

// ---- pass 3 --------------------------------------------------------------------------------------
// The walk of lane 0 over a trace's chunks with the true state, for traces in which not every chunk is
// fine. Chunk k's record was replayed from en(k) (pass 1: the guess after its warm-up; pass 2b and the
// second replay: the state they were given) and ended with `exit`; it stands if en(k) has the true
// future, or after a shift by the replica offset (tp_shift_safe). Otherwise:
//  RESOLVE (pass 3a, before the second replay): the chunk is marked (status 3) to be replayed a second
//    time from the true state, which is put where that replay reads it. What the chunk will end with is
//    not known yet, and the next chunk needs it: worked out for a chunk in which every tick met the same
//    threshold (tp_pure_exit), assumed to be what the first replay ended with otherwise (most chunks
//    forget the state they were entered with). Wrong assumptions cost time, not correctness:
//  FINAL (pass 3b, after the second replay): the same walk over what is there now. A chunk that still
//    does not fit is replayed here, up to kChainReplays per trace; a trace that needs more is flagged for
//    the one-trace-per-thread kernel that follows and nothing is written for it.

// One chunk of a trace as the walks need it. A walk is one thread going through a trace's chunks in
// order, and reading a chunk's status, states and record where it needs them costs it two memory latencies
// per chunk (which exit state counts depends on the status): 0.11 ms per walk at 4096 x 10^5, as much as a
// fifth of pass 1. So the warp's lanes fetch 32 chunks at a time into shared memory, each lane one chunk
// with all its loads in flight together, and lane 0 walks over those.
struct TpStaged {
  ksa_summary q;
  TpG used, exit;  // what the record was replayed from, and what it ended with (pass 2b's where the record is)
  int status;
};
__device__ __forceinline__ void tp_stage(const TpSumArgs& a, long long n, int k0, TpStaged* slots, int lane) {
  __syncwarp();  // (lane 0 is done with what is there)
  const int k = k0 + lane;
  if (k < a.K) {
    const long long idx = (long long)k * a.N + n;
    const int status = a.status[idx];
    const TpG io = tp_get(a.io, idx), fx = tp_get(a.fx, idx);  // both: neither load waits for the status
    slots[lane].q = a.part[idx];
    slots[lane].used = tp_get(a.en, idx);
    slots[lane].exit = (status == 1 || status < 0) ? fx : io;
    slots[lane].status = status;
  }
  __syncwarp();
}

// Chunks k0 .. k1 - 1 (staged in `slots`) of the walk; false: the trace is given up.
template <int A, bool FAULTS, bool RESOLVE>
__device__ bool tp_sum_walk(const TpSumArgs& a, const ksa_step_consts& c, long long n, int k0, int k1, const TpStaged* slots,
                            TpG& cur, TpRun& run, int& replays) {
  TpRun one;
  for (int k = k0; k < k1; ++k) {
    const long long idx = (long long)k * a.N + n;
    const int t0 = k * a.L;
    const int len = min(a.L, a.T - t0);
    const long long B = a.tick0 + t0;
    const bool last = k == a.K - 1;
    const TpG used = slots[k - k0].used;
    TpG exit = slots[k - k0].exit;
    ksa_summary q = slots[k - k0].q;
    const bool gates = tp_same_gates(a.pk, c, used, cur, B, a.T - t0);
    const long long delta = (long long)cur.r - used.r;
    if (gates && (delta == 0 || tp_shift_safe(c, q, used.r, delta))) {
      if (delta != 0) tp_shift(q, exit, delta, len, last ? a.gE_lo : a.gL_lo);
      tp_run_of_chunk(one, q, exit, B, len, false, last ? a.mE_lo : a.mL_lo, last ? a.mE_hi : a.mL_hi);
      cur.r = exit.r;
      cur.ju = tp_taken(exit.ju, B, len) ? exit.ju : cur.ju;
      cur.jd = tp_taken(exit.jd, B, len) ? exit.jd : cur.jd;
    } else if (RESOLVE) {
      const int kind = tp_pure_kind<FAULTS>(a.pk, q, len);
      TpG after = cur;
      if (kind != 0 && tp_pure_exit(a.pk, c, kind, B, len, after)) {
        // every tick met the same threshold: the state after the chunk is worked out here, its record from the
        // true entry state by the warp's lanes after the walk (status 7 / 8 = kind, then 6: tp_sum_resolve_kernel)
        tp_put(a.en, idx, cur);
        tp_put(a.io, idx, after);
        a.status[idx] = 6 + kind;
        cur = after;
        continue;
      }
      tp_put(a.en, idx, cur);
      tp_put(a.io, idx, cur);
      tp_summary_clear(q);
      a.part[idx] = q;
      a.status[idx] = 3;
      cur.r = exit.r;
      cur.ju = tp_taken(exit.ju, B, len) ? exit.ju : cur.ju;
      cur.jd = tp_taken(exit.jd, B, len) ? exit.jd : cur.jd;
      continue;  // (pass 3a builds no result)
    } else {
      if (++replays > kChainReplays) return false;  // left to the plain kernel that follows (g_tp_unresolved)
      tp_summary_clear(q);
      tp_sum_replay<A, FAULTS>(a, c, n, t0, len, cur, &q);
      tp_run_of_chunk(one, q, cur, B, len, true, last ? a.mE_lo : a.mL_lo, last ? a.mE_hi : a.mL_hi);
    }
    tp_run_append(run, one);
  }
  return true;
}
// The whole walk by a warp: lane 0 walks, all lanes stage. `run` and `unresolved` are lane 0's.
template <int A, bool FAULTS, bool RESOLVE>
__device__ void tp_sum_walk_warp(const TpSumArgs& a, const ksa_step_consts& c, long long n, TpStaged* slots, int lane,
                                 TpRun& run, bool& unresolved) {
  TpG cur = tp_get(a.io, n);  // chunk 0 ran from the caller's state
  int replays = 0;
  bool going = true;
  if (lane == 0) {
    tp_run_init(run);
    TpRun one;
    tp_run_of_chunk(one, a.part[n], cur, a.tick0, a.L, true, a.mL_lo, a.mL_hi);  // (K >= 2: chunk 0 has L ticks)
    tp_run_append(run, one);
  }
  for (int k0 = 1; k0 < a.K; k0 += 32) {
    tp_stage(a, n, k0, slots, lane);
    if (lane == 0 && going) going = tp_sum_walk<A, FAULTS, RESOLVE>(a, c, n, k0, min(k0 + 32, a.K), slots, cur, run, replays);
  }
  if (lane == 0) unresolved = !going;
  __syncwarp();
}

// ---- pass 3a: warp (trace) ----------------------------------------------------------------------
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_sum_resolve_kernel(const __grid_constant__ TpSumArgs a) {
  __shared__ TpStaged stage[BLOCK / 32][32];
  const long long n = (long long)blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
  if (n >= a.N) return;  // (whole warps)
  const int lane = threadIdx.x & 31;
  bool fine = true;
  for (int k = lane; k < a.K; k += 32) {
    const int status = a.status[(long long)k * a.N + n];
    fine = fine && (status == 0 || status == 1);
  }
  if (__all_sync(0xffffffffu, fine)) return;
  const ksa_step_consts c(a.pk);
  TpRun run;
  bool unresolved = false;
  tp_sum_walk_warp<A, FAULTS, true>(a, c, n, stage[threadIdx.x >> 5], lane, run, unresolved);
  // the records of the chunks the walk worked its way through in closed form: a few hundred instructions
  // each (powers and geometric sums), shared out over the lanes
  for (int k = 1 + lane; k < a.K; k += 32) {
    const long long idx = (long long)k * a.N + n;
    const int status = a.status[idx];
    if (status != 7 && status != 8) continue;
    const int len = min(a.L, a.T - k * a.L);
    const TpG entry = tp_get(a.en, idx);
    ksa_summary q;
    TpG after;
    if (tp_pure_record(a.pk, c, status - 6, a.tick0 + (long long)k * a.L, len, entry, q, after)) {
      a.part[idx] = q;
      a.status[idx] = 6;
    } else {  // (too many distinct counts to follow: replayed a second time after all, from the same entry state)
      tp_put(a.io, idx, entry);
      tp_summary_clear(q);
      a.part[idx] = q;
      a.status[idx] = 3;
    }
  }
}

// Does every chunk k0 .. k1 - 1 (staged) fit the true state now, as it stands or shifted by a replica offset?
// The walk without building the result. Leaves each chunk's offset in g_tp_delta (summary mode has no other
// use for it). False: some chunk does not fit (the full walk decides what then).
__device__ bool tp_sum_verify(const TpSumArgs& a, const ksa_step_consts& c, long long n, int k0, int k1, const TpStaged* slots,
                              TpG& cur) {
  for (int k = k0; k < k1; ++k) {
    const TpStaged& x = slots[k - k0];
    const int t0 = k * a.L;
    const int len = min(a.L, a.T - t0);
    const long long B = a.tick0 + t0;
    const long long delta = (long long)cur.r - x.used.r;
    if (!tp_same_gates(a.pk, c, x.used, cur, B, a.T - t0)) return false;
    if (delta != 0 && !tp_shift_safe(c, x.q, x.used.r, delta)) return false;
    if (delta != (long long)(int)delta) return false;  // (an offset of 2^31 or more: not one g_tp_delta holds)
    g_tp_delta[(long long)k * a.N + n] = (int)delta;
    cur.r = (int)(x.exit.r + delta);
    cur.ju = tp_taken(x.exit.ju, B, len) ? x.exit.ju : cur.ju;
    cur.jd = tp_taken(x.exit.jd, B, len) ? x.exit.jd : cur.jd;
  }
  return true;
}

// ---- pass 3b: warp (trace) ----------------------------------------------------------------------
template <int A, bool FAULTS>
__global__ void __launch_bounds__(BLOCK) tp_sum_chain_kernel(const __grid_constant__ TpSumArgs a) {
  __shared__ TpStaged stage[BLOCK / 32][32];
  const long long n = (long long)blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
  if (n >= a.N) return;  // (whole warps)
  const int lane = threadIdx.x & 31;
  TpStaged* const slots = stage[threadIdx.x >> 5];
  const ksa_step_consts c(a.pk);
  const int per = (a.K + 31) / 32;
  const int k0 = lane * per, k1 = min(a.K, k0 + per);
  TpRun run;
  bool fine = true, unresolved = false;
  // this lane's run of chunks, in order; `shifted`: each chunk by the offset tp_sum_verify left for it
  auto my_chunks = [&](const bool shifted) {
    tp_run_init(run);
    for (int k = k0; k < k1; ++k) {
      const long long idx = (long long)k * a.N + n;
      const int status = a.status[idx];
      fine = fine && (status == 0 || status == 1);
      const bool last = k == a.K - 1;
      const int len = min(a.L, a.T - k * a.L);
      ksa_summary q = a.part[idx];
      TpG exit = (status == 1 || status < 0) ? tp_get(a.fx, idx) : tp_get(a.io, idx);
      if (shifted && k > 0) {
        const int delta = g_tp_delta[idx];
        if (delta != 0) tp_shift(q, exit, delta, len, last ? a.gE_lo : a.gL_lo);
      }
      TpRun one;
      tp_run_of_chunk(one, q, exit, a.tick0 + (long long)k * a.L, len, k == 0, last ? a.mE_lo : a.mL_lo,
                      last ? a.mE_hi : a.mL_hi);
      tp_run_append(run, one);
    }
  };
  my_chunks(false);
  fine = __all_sync(0xffffffffu, fine);
  bool parallel = fine;
  if (!fine) {
    // not every chunk simply stands. If each fits the true state as it is or shifted (lane 0 looks), the
    // lanes combine them as above, with the offsets; else lane 0 walks and replays.
    bool fits = true;
    TpG cur = tp_get(a.io, n);  // chunk 0 ran from the caller's state
    for (int s0 = 1; s0 < a.K; s0 += 32) {
      tp_stage(a, n, s0, slots, lane);
      if (lane == 0 && fits) fits = tp_sum_verify(a, c, n, s0, min(s0 + 32, a.K), slots, cur);
    }
    parallel = __shfl_sync(0xffffffffu, (int)fits, 0) != 0;
    __syncwarp();  // (lane 0's offsets are the other lanes' to read)
    if (parallel) my_chunks(true);
  }
  if (parallel) {
    // every chunk was replayed from (the equivalent of) the true state: combine the lanes' runs in order
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      TpRun other;
      tp_run_shfl(other, run, min(lane + off, 31));
      if ((lane & (2 * off - 1)) == 0) tp_run_append(run, other);
    }
  } else {
    tp_sum_walk_warp<A, FAULTS, false>(a, c, n, slots, lane, run, unresolved);
  }
  if (lane != 0) return;
  g_tp_unresolved[n] = unresolved ? 1 : 0;
  if (unresolved) return;
  // add the call to the caller's record (a read-modify-write: calls accumulate) and write the state back
  ksa_summary q = a.summary[n];
  q.final_replicas = run.r;
  q.min_replicas = min(q.min_replicas, run.rmin);
  q.max_replicas = max(q.max_replicas, run.rmax);
  q.up_applied += run.cnt[0];
  q.down_applied += run.cnt[1];
  q.up_noop += run.cnt[2];
  q.down_noop += run.cnt[3];
  q.up_cool += run.cnt[4];
  q.down_cool += run.cnt[5];
  q.poll_failed += run.cnt[6];
  q.up_failed += run.cnt[7];
  q.down_failed += run.cnt[8];
  q.sum_replicas = (long long)((unsigned long long)q.sum_replicas + (unsigned long long)run.sum);
  const unsigned lo = (unsigned)q.hash * run.m_lo + run.h_lo;
  const unsigned hi = (unsigned)(q.hash >> 32) * run.m_hi + run.h_hi;
  q.hash = ((unsigned long long)hi << 32) | lo;
  a.summary[n] = q;
  a.st_r[n] = run.r;
  a.st_ju[n] = run.ju;  // (chunk 0 always carries a stamp: has_u / has_d are set)
  a.st_jd[n] = run.jd;
}

// One time-parallel replay per device at a time (the scratch is shared): `busy` is recorded after
// the last launch of a call and queried by the next one.
constexpr int kMaxDevices = 64;
std::mutex g_tp_mutex;
cudaEvent_t g_tp_busy[kMaxDevices] = {};
cudaStream_t g_tp_stream[kMaxDevices] = {};  // stream of the call that recorded `busy`
// Device addresses of the scratch arrays: in host code their names are only stand-ins for them
// (cudaGetSymbolAddress, once per device, with that device current).
char* g_tp_scratch_addr[kMaxDevices] = {};
ksa_summary* g_tp_partial_addr[kMaxDevices] = {};
int* g_tp_unresolved_addr[kMaxDevices] = {};

int tp_addresses(int device) {
  if (g_tp_scratch_addr[device]) return KSA_OK;
  void* p = nullptr;
  void* q = nullptr;
  void* u = nullptr;
  KSA_CUDA(cudaGetSymbolAddress(&p, g_tp_scratch));
  KSA_CUDA(cudaGetSymbolAddress(&q, g_tp_partial));
  KSA_CUDA(cudaGetSymbolAddress(&u, g_tp_unresolved));

  g_tp_partial_addr[device] = (ksa_summary*)q;
  g_tp_unresolved_addr[device] = (int*)u;
  g_tp_scratch_addr[device] = (char*)p;
  return KSA_OK;
}

// Takes the scratch of the current device for a call on `st`. Returns KSA_OK with the device index, an
// error, or 1: the scratch is in use by an earlier call on another stream, or `st` is being captured.
// The caller holds g_tp_mutex and calls tp_release after its last launch.
int tp_acquire(cudaStream_t st, int* device_out) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) {
    cudaGetLastError();
    return 1;
  }
  if (cap != cudaStreamCaptureStatusNone) return 1;
  int device = -1;
  KSA_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= kMaxDevices) return 1;
  if (!g_tp_busy[device]) {
    KSA_CUDA(cudaEventCreateWithFlags(&g_tp_busy[device], cudaEventDisableTiming));
  } else if (cudaEventQuery(g_tp_busy[device]) != cudaSuccess) {
    cudaGetLastError();
    // On the same stream the new call is ordered behind the old one anyway; on another stream the
    // two would overlap on the scratch: plain path.
    if (st != g_tp_stream[device]) return 1;
  }
  *device_out = device;
  return KSA_OK;
}
int tp_release(cudaStream_t st, int device) {
  KSA_CUDA(cudaEventRecord(g_tp_busy[device], st));
  g_tp_stream[device] = st;
  return KSA_OK;
}

template <int A, bool FAULTS>
int tp_launch(const TpArgs& a, bool wide, cudaStream_t st) {
  if (a.probe_limit != INT_MAX) {
    const long long pairs = (long long)a.probe_traces * (a.K - 1);
    tp_probe_kernel<A, FAULTS><<<(unsigned)((pairs + 31) / 32), 96, 0, st>>>(a);
    KSA_LAUNCH_CHECK("tp_probe_kernel");
  }
  if (wide) {
    if constexpr (FAULTS) {
      // 4 traces per thread, the shape of ksa_replay_full_faults' default: with eight interleaved the
      // failure-aware step runs out of issue slots (1.68 ms against 0.94 without a mask at 4096 x 10^5)
      const unsigned gx = (unsigned)((a.N / 4 + BLOCK - 1) / BLOCK);
      tp_spec_kernel<A, 4, A == 1 ? 4 : (A == 2 ? 2 : 1), BLOCK, A == 3 ? 4 : 5, true><<<dim3(gx, (unsigned)a.K), BLOCK, 0, st>>>(a);
    } else {  // 8 traces per thread, the launch shape of the main kernel's default
      // Blocks of 128 threads where the grid is a fraction of a wave (4096 x 10^5: 96 chunks x 2 blocks of 256 on
      // 296 slots leave SMs with one block next to SMs with two; 4 blocks of 128: 1.121 -> 1.080 ms, 16384 x 40000:
      // 1.623 -> 1.574; blocks of 64 the same). Above the limit (the second wave: two chunks, many blocks) 256.
      if (a.N <= kTpMaxFull) {
        constexpr int TPB = 128;
        const unsigned gx = (unsigned)((a.N / 8 + TPB - 1) / TPB);
        tp_spec_kernel<A, 8, A == 1 ? 2 : 1, TPB, A == 1 ? 4 : 2, false><<<dim3(gx, (unsigned)a.K), TPB, 0, st>>>(a);
      } else {
        constexpr int TPB = 256;
        const unsigned gx = (unsigned)((a.N / 8 + TPB - 1) / TPB);
        tp_spec_kernel<A, 8, A == 1 ? 2 : 1, TPB, A == 1 ? 2 : 1, false><<<dim3(gx, (unsigned)a.K), TPB, 0, st>>>(a);
      }
    }
  } else {
    const unsigned gx = (unsigned)((a.N + BLOCK - 1) / BLOCK);
    tp_spec_kernel<A, 1, 8, BLOCK, A == 1 ? 8 : (A == 2 ? 5 : 4), FAULTS><<<dim3(gx, (unsigned)a.K), BLOCK, 0, st>>>(a);
  }
  KSA_LAUNCH_CHECK("tp_spec_kernel");
  const unsigned gn = (unsigned)((a.N + BLOCK - 1) / BLOCK);
  tp_check_kernel<<<dim3(gn, (unsigned)(a.K - 1)), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_check_kernel");
  const long long items = a.N * (a.K - 1);  // (the list's length is only known on the device: idle threads return at once)
  tp_fix_kernel<A, FAULTS><<<(unsigned)((items + BLOCK - 1) / BLOCK), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_fix_kernel");
  // traces with chunks that neither stand nor got patched: what kind of chunks they have, their true entry
  // states, a second replay of the chunks that need one (threads of the other chunks return at once)
  tp_resolve_kernel<A, FAULTS><<<gn, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_resolve_kernel");
  tp_spec_kernel<A, 1, 8, BLOCK, A == 1 ? 8 : (A == 2 ? 5 : 4), FAULTS, true><<<dim3(gn, (unsigned)(a.K - 1)), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_spec_kernel (second replay)");
  if constexpr (!FAULTS) {  // (with a failure mask no chunk has a closed form)
    tp_fill_kernel<<<dim3((unsigned)((a.N + 8 * BLOCK - 1) / (8 * BLOCK)), (unsigned)(a.K - 1), kFillParts), BLOCK, 0, st>>>(a);
    KSA_LAUNCH_CHECK("tp_fill_kernel");
  }
  tp_chain_kernel<A, FAULTS><<<gn, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_chain_kernel");
  tp_shift_fill_kernel<<<dim3((unsigned)((a.N + 8 * BLOCK - 1) / (8 * BLOCK)), (unsigned)(a.K - 1), kFillParts), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_shift_fill_kernel");
  tp_shift_kernel<<<dim3(gn, (unsigned)(a.K - 1)), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_shift_kernel");
  return KSA_OK;
}

template <int A, bool FAULTS>
int tp_sum_launch(const TpSumArgs& a, bool ordered, cudaStream_t st) {
  const unsigned gn = (unsigned)((a.N + BLOCK - 1) / BLOCK);
  tp_sum_warm_kernel<A, FAULTS><<<dim3(gn, (unsigned)a.K), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_sum_warm_kernel");
  const int rc = ksa_replay_summary_chunks(a.pk, ordered, (const int64_t*)a.depth, a.fail, a.L, a.T, A, a.N, a.ld, a.tick0,
                                           a.K, a.io.r, (int64_t*)a.io.ju, (int64_t*)a.io.jd, a.part, nullptr, 0, st);
  if (rc != KSA_OK) return rc;
  tp_sum_check_kernel<FAULTS><<<dim3(gn, (unsigned)(a.K - 1)), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_sum_check_kernel");
  // (the list's length is only known on the device: a warp per possible item, idle warps return at once)
  const long long items = a.N * (a.K - 1);
  const long long warps_per_block = BLOCK / 32;
  tp_sum_fix_kernel<A, FAULTS><<<(unsigned)((items + warps_per_block - 1) / warps_per_block), BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_sum_fix_kernel");
  const unsigned per_trace = (unsigned)((a.N + warps_per_block - 1) / warps_per_block);
  tp_sum_resolve_kernel<A, FAULTS><<<per_trace, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_sum_resolve_kernel");
  // the chunks pass 3a marked, a second time, from the entry states it put into `io` (the same grid as pass 1,
  // so that neighbouring traces still share their loads; the other threads return at once)
  const int rc2 = ksa_replay_summary_chunks(a.pk, ordered, (const int64_t*)a.depth, a.fail, a.L, a.T, A, a.N, a.ld, a.tick0,
                                            a.K, a.io.r, (int64_t*)a.io.ju, (int64_t*)a.io.jd, a.part, a.status, 3, st);
  if (rc2 != KSA_OK) return rc2;
  tp_sum_chain_kernel<A, FAULTS><<<per_trace, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("tp_sum_chain_kernel");
  return KSA_OK;
}

unsigned pow_u32(unsigned base, unsigned e) {
  unsigned r = 1u;
  for (; e; e >>= 1, base *= base)
    if (e & 1u) r *= base;
  return r;
}
// 1 + m + m^2 + ... + m^(n-1) mod 2^32, by halving: G(2k) = G(k) (1 + m^k), G(2k + 1) = G(2k) m + 1
unsigned geo_u32(unsigned m, unsigned n) {
  if (n == 0) return 0u;
  if (n & 1u) return geo_u32(m, n - 1) * m + 1u;
  return geo_u32(m, n / 2) * (1u + pow_u32(m, n / 2));
}

}  // namespace

// How the time-parallel replay would cut this call: chunks K (0 = take the plain kernels: too many
// traces to gain, too few ticks, or no room in the scratch), ticks per chunk L, warm-up ticks, and
// whether pass 1 can run several traces per thread (`vector_ok`: full mode with the layout the vector
// kernels need and N a multiple of 8; summary mode runs one trace per thread). Environment, for
// tests and measurements: KSA_TIMEPAR_CHUNKS=<K> forces K chunks whenever that is possible at all
// (0 turns the path off), KSA_TIMEPAR_WARM=<ticks> the warm-up, KSA_TIMEPAR_WIDE=0 one trace per thread.
int ksa_timepar_plan(int64_t T, int64_t N, bool summary, bool faults, bool vector_ok, int* L_out, int* warm_out,
                     int* wide_out) {
  int forced = -1;
  if (const char* e = std::getenv("KSA_TIMEPAR_CHUNKS")) forced = std::atoi(e);
  if (forced == 0) return 0;
  bool wide = vector_ok && !summary;
  if (const char* e = std::getenv("KSA_TIMEPAR_WIDE")) wide = wide && std::atoi(e) != 0;
  int warm = 128;  // profiles/r02_small_n.txt: 0.80 of the copy peak at 4096 x 10^5 against 0.78 with 64, 0.68 with none
  if (const char* e = std::getenv("KSA_TIMEPAR_WARM")) warm = std::max(0, std::atoi(e)) / 8 * 8;
  long long K;
  if (forced > 0) {
    K = forced;
    if (K < 2 || (T + K - 1) / K < 16) return 0;
  } else {
    if (T < 4096) return 0;
    const long long kSMs = 148;
    if (summary || (faults && wide)) {
      // Pass 1 runs blocks of 128 threads (one trace each; with a mask in full mode four each), 8 per SM
      // (7 / 5 with a failure mask): K is the largest count whose grid fills whole waves of those
      // slots: a grid of 2.3 waves takes as long as one of 3 (profiles/r02_small_n_modes.txt).
      const long long per_block = summary ? 128 : 512;
      const long long slots = kSMs * (summary ? (faults ? 7 : 8) : 5);
      const long long waves = summary ? 2 : 1;
      const long long blocks_x = (N + per_block - 1) / per_block;
      if (summary ? N > (faults ? kTpMaxSummaryFaults : kTpMaxSummary)
                  : (N > kTpMaxFullFaults && N <= kTpFullFaultsSecondWave))
        return 0;
      K = waves * slots / blocks_x;
    } else {
      // One trace per thread fills the GPU at about 2^18 threads, eight per thread (1-KB requests, the
      // shape that reaches the copy peak) at about 400k trace-chunks (profiles/r01_small_n_variants.txt).
      if (N > kMaxTraces || (N > kTpMaxFull && N <= kTpFullSecondWave)) return 0;
      const long long want = wide ? 393216 : 262144;
      K = (want + N - 1) / N;
    }
    // Chunks of at least 8 x warm ticks keep the warm-up and the re-replayed prefixes small.
    const long long kmax = T / (8 * std::max(warm, 64));
    if (K > kmax) K = kmax;
    if (K > 512) K = 512;
    if (K < 2) return 0;
  }
  long long L = ((T + K - 1) / K + 7) / 8 * 8;
  K = (T + L - 1) / L;  // (rounding L up may leave fewer chunks)
  if (K < 2 || N * K > kMaxChunks || N > kMaxTraces) return 0;
  if (warm > L) warm = (int)L;
  *L_out = (int)L;
  *warm_out = warm;
  *wide_out = wide ? 1 : 0;
  return (int)K;
}

// ksa_replay_full / ksa_replay_full_faults (fail_dev != NULL) for (T, N) with both outputs, in K
// chunks per trace. Returns KSA_OK, an error, or 1 if the scratch is in use by an earlier call on
// another stream of this device (or the stream is being captured): the caller then takes the plain path.
int ksa_timepar_replay_full(const ksa_policy_k& pk, const int64_t* depth_dev, const uint8_t* fail_dev, int64_t T,
                            int32_t A, int64_t N, int64_t ld, int64_t tick0, const ksa_state* state,
                            int32_t* replicas_dev, uint8_t* events_dev, int K, int L, int warm, int wide,
                            const ksa_plain_launch& plain, cudaStream_t st) {
  std::lock_guard<std::mutex> lock(g_tp_mutex);
  int device = -1;
  int rc = tp_acquire(st, &device);
  if (rc != KSA_OK) return rc;
  rc = tp_addresses(device);
  if (rc != KSA_OK) return rc;
  TpArgs a;
  a.pk = pk;
  a.depth = (const long long*)depth_dev;
  a.fail = fail_dev;
  a.ld = ld;
  a.tick0 = tick0;
  a.N = N;
  a.T = (int)T;
  a.A = A;
  a.K = K;
  a.L = L;
  a.warm = warm;
  a.fix_limit = (int)std::max<int64_t>(64, N * (K - 1) / 32);
  // Up to 16384 traces even two chunked replays beat the plain kernel; beyond, a probe decides first
  // (KSA_TIMEPAR_PROBE=0 / 1 forces it off / on). With a failure mask it always does: saturated / idle
  // stretches have no closed form there, and a trace that stays unresolved costs the plain kernel's whole
  // latency on top (4096 x 10^5 such queues: 18.1 ms against 10.3).
  bool probe = N > 16384 || fail_dev != nullptr;
  if (const char* e = std::getenv("KSA_TIMEPAR_PROBE")) probe = std::atoi(e) != 0;
  // about 8192 (trace, chunk start) pairs: the probe's time is its longest warm-up, not its size, as long as it is small
  a.probe_traces = (int)std::min<int64_t>(N / 32 * 32, std::min<int64_t>(512, std::max<int64_t>(32, 8192 / (K - 1) / 32 * 32)));
  a.probe_stride = a.probe_traces >= 32 ? (int)(N / (a.probe_traces / 32)) : 0;
  if (a.probe_traces < 32) probe = false;
  a.probe_limit = probe ? (int)((int64_t)a.probe_traces * (K - 1) / 8) : INT_MAX;
  a.one = 1;
  a.st_r = state->replicas_dev;
  a.st_ju = (long long*)state->last_up_tick_dev;
  a.st_jd = (long long*)state->last_down_tick_dev;
  a.replicas = replicas_dev;
  a.events = events_dev;
  const bool w = wide != 0;
  if (fail_dev)
    rc = A == 1 ? tp_launch<1, true>(a, w, st) : A == 2 ? tp_launch<2, true>(a, w, st) : tp_launch<3, true>(a, w, st);
  else
    rc = A == 1 ? tp_launch<1, false>(a, w, st) : A == 2 ? tp_launch<2, false>(a, w, st) : tp_launch<3, false>(a, w, st);
  if (rc != KSA_OK) return rc;
  rc = plain(g_tp_unresolved_addr[device]);  // the traces pass 3 gave up, one thread each, from the caller's state
  if (rc != KSA_OK) return rc;
  return tp_release(st, device);
}

// ksa_replay_summary / ksa_replay_summary_faults (fail_dev != NULL) in K chunks per trace; returns as above.
int ksa_timepar_replay_summary(const ksa_policy_k& pk, bool ordered, const int64_t* depth_dev, const uint8_t* fail_dev,
                               int64_t T, int32_t A, int64_t N, int64_t ld, int64_t tick0, const ksa_state* state,
                               ksa_summary* summary_dev, int K, int L, int warm, const ksa_plain_launch& plain,
                               cudaStream_t st) {
  std::lock_guard<std::mutex> lock(g_tp_mutex);
  int device = -1;
  int rc = tp_acquire(st, &device);
  if (rc != KSA_OK) return rc;
  rc = tp_addresses(device);
  if (rc != KSA_OK) return rc;
  TpSumArgs a;
  a.pk = pk;
  a.depth = (const long long*)depth_dev;
  a.fail = fail_dev;
  a.ld = ld;
  a.tick0 = tick0;
  a.N = N;
  a.T = (int)T;
  a.A = A;
  a.K = K;
  a.L = L;
  a.warm = warm;
  a.one = 1;
  a.piece = std::max(16, ((L + 31) / 32 + 7) / 8 * 8);
  a.fix_limit = (int)std::max<int64_t>(64, N * (K - 1) / 32);
  const int lenE = (int)(T - (int64_t)(K - 1) * L);
  a.mL_lo = pow_u32(KSA_HASH_MUL_LO, (unsigned)L);
  a.mL_hi = pow_u32(KSA_HASH_MUL_HI, (unsigned)L);
  a.mE_lo = pow_u32(KSA_HASH_MUL_LO, (unsigned)lenE);
  a.mE_hi = pow_u32(KSA_HASH_MUL_HI, (unsigned)lenE);
  a.gL_lo = geo_u32(KSA_HASH_MUL_LO, (unsigned)L);
  a.gE_lo = geo_u32(KSA_HASH_MUL_LO, (unsigned)lenE);
  a.st_r = state->replicas_dev;
  a.st_ju = (long long*)state->last_up_tick_dev;
  a.st_jd = (long long*)state->last_down_tick_dev;
  a.summary = summary_dev;
  // the 64 bytes per slot of g_tp_scratch, as arrays over the kMaxChunks slots
  char* base = g_tp_scratch_addr[device];
  const size_t M = (size_t)kMaxChunks;
  auto carry = [&](size_t off) {
    TpCarry c;
    c.r = (int*)(base + off * M);
    c.ju = (long long*)(base + (off + 4) * M);
    c.jd = (long long*)(base + (off + 12) * M);
    return c;
  };
  a.io = carry(0);
  a.en = carry(20);
  a.fx = carry(40);
  a.status = (int*)(base + 60 * M);
  a.part = g_tp_partial_addr[device];

  if (fail_dev)
    rc = A == 1 ? tp_sum_launch<1, true>(a, ordered, st)
                : A == 2 ? tp_sum_launch<2, true>(a, ordered, st) : tp_sum_launch<3, true>(a, ordered, st);
  else
    rc = A == 1 ? tp_sum_launch<1, false>(a, ordered, st)
                : A == 2 ? tp_sum_launch<2, false>(a, ordered, st) : tp_sum_launch<3, false>(a, ordered, st);
  if (rc != KSA_OK) return rc;
  rc = plain(g_tp_unresolved_addr[device]);
  if (rc != KSA_OK) return rc;
  return tp_release(st, device);
}
