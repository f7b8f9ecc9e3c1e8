// ksa_replay.cu — the hot path: ksa_replay_full and ksa_replay_summary.
//
// One thread owns V adjacent traces and walks down the T ticks of the tick-major [T][A][ld] int64
// depth array. Per tick it issues one 128/256-bit coalesced load per attribute plane (LDG.E.128 /
// LDG.E.256 on sm_100a), D ticks ahead of use (a register ring, so D*V*A*8 bytes per thread are
// in flight and the loop-carried state chain never waits on DRAM), runs the branch-free policy
// step (ksa_policy.cuh) on registers, and writes the per-tick replica counts (int32) and event
// bytes back as one vector store each. Nothing is re-read, so all accesses carry streaming hints
// (L1::no_allocate, evict-first). No shared memory: there is no reuse to stage (a cp.async.bulk /
// mbarrier staging of the rows exists for the summary replay, ksa_summary_bulk_kernel below; it
// measured slower than the register ring and is kept as a tuning point).
//
// Replaces, for N traces at once: sqs/sqs.go:56-64 (sum), main.go:51-77 (thresholds, cool-downs,
// stamps), scale/scale.go:60-69,87-97 (clamped steps). HBM-bound: 8A bytes in + 5 bytes out per
// (trace, tick) in full mode, 8A in summary mode.
#include "ksa_internal.h"

#include <cstdlib>

namespace {

struct ReplayArgs {
  ksa_policy_k pk;
  const long long* depth;
  const unsigned char* fail;  // *_faults entry points: [T][ld] KSA_FAIL_* bits, else NULL
  long long ld;      // row pitch, elements
  long long tick0;
  long long n_begin; // first trace of this launch
  long long n_count; // traces of this launch (a multiple of V)
  int T;
  int one;  // always 1: a multiplier ptxas cannot fold (ksa_step_fma)
  int* st_r;
  long long* st_ju;
  long long* st_jd;
  int* replicas;           // full mode, may be NULL
  unsigned char* events;   // full mode, may be NULL
  ksa_summary* summary;    // summary mode
  // Time-parallel summary launches only (TP kernels, ksa_timepar.cu): the call's T ticks are cut
  // into chunks of `T` ticks (above), blockIdx.y = chunk, and chunk k of trace n keeps its carry
  // state and its summary at index k * tp_stride + n of the arrays above (scratch, not the caller's).
  int tp_total;            // ticks of the whole call
  long long tp_stride;     // traces of the call
  // KSA_GATED kernels only: replay trace n only if gate[n] != 0. The launch that follows the chunked
  // passes of ksa_timepar.cu, for the traces those gave up. KSA_CHUNKS_MARKED kernels: replay chunk k of
  // trace n only if gate[k * tp_stride + n] == gate_value.
  const int* gate;
  int gate_value;
};

// MODE_FULL writes both outputs; MODE_FULL_OPT tests the two output pointers at run time (either may
// be NULL); MODE_SUMMARY writes nothing per tick.
enum { MODE_FULL = 0, MODE_SUMMARY = 1, MODE_FULL_OPT = 2 };
// STEP: 0 = ksa_step (C++), 1 = ksa_step_ptx (predication spelled out in PTX), 2 = ksa_step_fma
// (conditional moves/adds as predicated IMADs on the FMA pipe). Same function, same results.
// 4 = ksa_step_faults (C++), 5 = ksa_step_faults_fma / ksa_step_faults_fma2: the steps of the
// *_faults entry points, which also read one failure-mask byte per decision.
// 6 = ksa_step_fma for ordered thresholds (summary mode), 7 = the same with failures,
// 8 = ksa_step_faults_fma2 with the event byte out (full mode), 9 / 10 = ksa_step_faults_fma3
// (summary mode; 9 for ordered thresholds). In summary mode the FMA steps
// return the wide event code (ksa_policy.cuh) and one trace per thread is the default.
// Load hints: 0 = ld.global.nc, 1 = + L1::no_allocate, 2 = + L2::evict_first (256-bit loads only)
enum { HINT_NC = 0, HINT_NA = 1, HINT_NA_EF = 2 };

// ---- vector loads / stores (inline PTX: 256-bit global accesses have no C++ spelling) ----------
template <int HINT>
__device__ __forceinline__ void ld64x4(const long long* p, long long& a, long long& b, long long& c,
                                       long long& d) {
  if (HINT == HINT_NA_EF)
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
  else if (HINT == HINT_NA)
    asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
  else
    asm volatile("ld.global.nc.v4.b64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
}
template <int HINT>
__device__ __forceinline__ void ld64x2(const long long* p, long long& a, long long& b) {
  if (HINT != HINT_NC)
    asm volatile("ld.global.nc.L1::no_allocate.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
  else
    asm volatile("ld.global.nc.v2.b64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
}
template <int HINT>
__device__ __forceinline__ void ld64x1(const long long* p, long long& a) {
  if (HINT != HINT_NC)
    asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(a) : "l"(p));
  else
    asm volatile("ld.global.nc.b64 %0, [%1];" : "=l"(a) : "l"(p));
}

template <int V, int HINT>
__device__ __forceinline__ void load_depth(const long long* p, long long (&v)[V]) {
  if constexpr (V == 1) {
    ld64x1<HINT>(p, v[0]);
  } else if constexpr (V == 2) {
    ld64x2<HINT>(p, v[0], v[1]);
  } else {
#pragma unroll
    for (int q = 0; q < V / 4; ++q) ld64x4<HINT>(p + 4 * q, v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
  }
}

template <int V>
__device__ __forceinline__ void store_replicas(int* p, const int (&r)[V]) {
  if constexpr (V == 1) {
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p), "r"(r[0]) : "memory");
  } else if constexpr (V == 2) {
    asm volatile("st.global.cs.v2.b32 [%0], {%1,%2};" ::"l"(p), "r"(r[0]), "r"(r[1]) : "memory");
  } else if constexpr (V == 4) {
    asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]) : "memory");
  } else {
    static_assert(V == 8 || V == 16, "V must be 1, 2, 4, 8 or 16");
#pragma unroll
    for (int q = 0; q < V / 8; ++q)
      asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p + 8 * q),
                   "r"(r[8 * q]), "r"(r[8 * q + 1]), "r"(r[8 * q + 2]), "r"(r[8 * q + 3]),
                   "r"(r[8 * q + 4]), "r"(r[8 * q + 5]), "r"(r[8 * q + 6]), "r"(r[8 * q + 7])
                   : "memory");
  }
}

// Pack V event bytes into words. FMA = true packs with multiply-adds (e1 * 256 + e0 ...), which
// ptxas emits as IMAD on the FMA pipe instead of shifts and LOP3 on the ALU pipe.
template <bool FMA>
__device__ __forceinline__ unsigned pack4(unsigned e0, unsigned e1, unsigned e2, unsigned e3) {
  if constexpr (FMA) {
    unsigned w;
    asm("mad.lo.u32 %0, %2, 256, %1;\n\t"
        "mad.lo.u32 %0, %3, 65536, %0;\n\t"
        "mad.lo.u32 %0, %4, 16777216, %0;"
        : "=&r"(w) : "r"(e0), "r"(e1), "r"(e2), "r"(e3));
    return w;
  } else {
    return e0 | (e1 << 8) | (e2 << 16) | (e3 << 24);
  }
}

template <int V, bool FMA>
__device__ __forceinline__ void store_events(unsigned char* p, const unsigned (&e)[V]) {
  if constexpr (V == 1) {
    asm volatile("st.global.cs.b8 [%0], %1;" ::"l"(p), "r"(e[0]) : "memory");
  } else if constexpr (V == 2) {
    const unsigned short w = (unsigned short)(e[0] | (e[1] << 8));
    asm volatile("st.global.cs.b16 [%0], %1;" ::"l"(p), "h"(w) : "memory");
  } else if constexpr (V == 4) {
    const unsigned w = pack4<FMA>(e[0], e[1], e[2], e[3]);
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p), "r"(w) : "memory");
  } else if constexpr (V == 8) {
    const unsigned w0 = pack4<FMA>(e[0], e[1], e[2], e[3]);
    const unsigned w1 = pack4<FMA>(e[4], e[5], e[6], e[7]);
    asm volatile("st.global.cs.v2.b32 [%0], {%1,%2};" ::"l"(p), "r"(w0), "r"(w1) : "memory");
  } else {
    static_assert(V == 16, "V must be 1, 2, 4, 8 or 16");
    const unsigned w0 = pack4<FMA>(e[0], e[1], e[2], e[3]);
    const unsigned w1 = pack4<FMA>(e[4], e[5], e[6], e[7]);
    const unsigned w2 = pack4<FMA>(e[8], e[9], e[10], e[11]);
    const unsigned w3 = pack4<FMA>(e[12], e[13], e[14], e[15]);
    asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
  }
}

// V failure-mask bytes of one tick, as packed words (same streaming hints as the depth loads).
template <int V>
__device__ __forceinline__ void load_fail(const unsigned char* p, unsigned (&w)[(V + 3) / 4]) {
  if constexpr (V == 1) {
    asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(w[0]) : "l"(p));
  } else if constexpr (V == 2) {
    unsigned short h;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(h) : "l"(p));
    w[0] = h;
  } else if constexpr (V == 4) {
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w[0]) : "l"(p));
  } else if constexpr (V == 8) {
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p));
  } else {
    static_assert(V == 16, "V must be 1, 2, 4, 8 or 16");
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
  }
}

typedef ksa_summary_acc SummaryRegs;

// LAUNCH: 0 = an ordinary call; KSA_CHUNKS = chunks of the time axis (TP below); KSA_GATED = only the
// traces flagged in a.gate (both for ksa_timepar.cu; separate instantiations, so that the kernels of
// ordinary calls are compiled exactly as before).
// KSA_CHUNKS_MARKED = KSA_CHUNKS for the (chunk, trace) pairs with a.gate[chunk * traces + trace] == a.gate_value only.
enum { KSA_CHUNKS = 1, KSA_GATED = 2, KSA_CHUNKS_MARKED = 3 };
template <int A, int V, int D, int HINT, int MODE, int STEP, int BLOCK, int MINB, int LAUNCH = 0>
__global__ void __launch_bounds__(BLOCK, MINB) ksa_replay_kernel(const __grid_constant__ ReplayArgs a) {
  const long long g = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (g * V >= a.n_count) return;
  const long long n0 = a.n_begin + g * V;
  constexpr bool TP = LAUNCH == KSA_CHUNKS || LAUNCH == KSA_CHUNKS_MARKED;
  if constexpr (LAUNCH == KSA_GATED) {
    static_assert(V == 1, "gated launches run one trace per thread");
    if (a.gate[n0] == 0) return;
  }
  ksa_step_consts c(a.pk);
  // TP: this block replays chunk blockIdx.y of the time axis as if it were a call of its own
  // (ticks tick0 + 1 .. tick0 + T of the trace, state and summary in the chunk's own slots).
  const int ky = TP ? (int)blockIdx.y : 0;
  const long long si = TP ? ky * a.tp_stride + n0 : n0;  // index into the state and summary arrays
  if constexpr (LAUNCH == KSA_CHUNKS_MARKED) {
    static_assert(V == 1, "marked chunks run one trace per thread");
    if (a.gate[si] != a.gate_value) return;
  }
  const int T = TP ? min(a.T, a.tp_total - ky * a.T) : a.T;
  const long long tick0 = TP ? a.tick0 + (long long)ky * a.T : a.tick0;
  const int one = a.one;  // always 1; see ksa_step_fma
  constexpr bool FAULTS = (STEP == 4 || STEP == 5 || STEP == 7 || STEP == 8 || STEP == 9 || STEP == 10);
  constexpr bool FMA = (STEP == 2 || STEP == 5 || STEP == 6 || STEP == 7 || STEP == 8 || STEP == 9 || STEP == 10);
  constexpr bool FAULTS3 = (STEP == 9 || STEP == 10);  // ksa_step_faults_fma3 counts the failures itself
  // Summary mode numbers the ticks of a group of D from 1 (compile-time constants after unrolling)
  // and moves the two stamps back by D after every group instead (see the loops below).
  constexpr bool REBASE = (MODE == MODE_SUMMARY);
  if constexpr (MODE == MODE_SUMMARY && V == 1) {
    // The pod steps and cool-downs are launch constants, which ptxas keeps in uniform registers;
    // the add of VIADDMNMX and the addend of the stamp's IMAD want them in ordinary ones and get a
    // copy per tick (3-4 of ~36 instructions per decision with one trace per thread). Adding a zero
    // that depends on the thread index ((one - 1) * threadIdx.x) makes them per-thread values.
    const int zero = (one - 1) * (int)threadIdx.x;
    c.up32 += zero;
    c.down32 += zero;
    if constexpr (FAULTS3) {  // (measured on SASS: 47.9 -> 47.6 per decision here, 32.3 -> 33.0 without failures)
      c.cu += zero;
      c.cd += zero;
    }
  }

  ksa_trace_regs s[V];
#pragma unroll
  for (int v = 0; v < V; ++v) {
    s[v].r = a.st_r[si + v];
    s[v].su = ksa_local_stamp(a.pk.cu, c.cu, tick0, a.st_ju[si + v]);
    s[v].sd = ksa_local_stamp(a.pk.cd, c.cd, tick0, a.st_jd[si + v]);
  }
  SummaryRegs sm[MODE == MODE_SUMMARY ? V : 1];
  if constexpr (MODE == MODE_SUMMARY) {
#pragma unroll
    for (int v = 0; v < V; ++v) {
      const ksa_summary* q = a.summary + si + v;
      ksa_summary_acc_init(sm[v], q->min_replicas, q->max_replicas, q->sum_replicas, q->hash);
    }
  }
  auto fold_counters = [&]() {
    if constexpr (MODE == MODE_SUMMARY) {
#pragma unroll
      for (int v = 0; v < V; ++v) ksa_summary_fold(sm[v]);
    }
  };
  auto flush_counters = [&]() {
    if constexpr (MODE == MODE_SUMMARY) {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        ksa_summary_flush(sm[v], a.summary + si + v);
        if constexpr (FAULTS3) ksa_summary_flush_faults3(sm[v], a.summary + si + v);
        else if constexpr (FAULTS) ksa_summary_flush_faults(sm[v], a.summary + si + v);
      }
    }
  };

  const long long row = (long long)A * a.ld;  // elements between consecutive ticks
  const long long* dp = a.depth + n0 + (TP ? (long long)ky * a.T * row : 0ll);
  int* rp = (MODE != MODE_SUMMARY) ? a.replicas + n0 : nullptr;
  unsigned char* ep = (MODE != MODE_SUMMARY) ? a.events + n0 : nullptr;
  const bool want_r = MODE == MODE_FULL || (MODE == MODE_FULL_OPT && a.replicas != nullptr);
  const bool want_e = MODE == MODE_FULL || (MODE == MODE_FULL_OPT && a.events != nullptr);

  // Register ring: buf[j] holds tick (t + j) for the current group of D ticks.
  long long buf[D][A][V];
#pragma unroll
  for (int j = 0; j < D; ++j)
    if (j < T) {
#pragma unroll
      for (int p = 0; p < A; ++p) load_depth<V, HINT>(dp + (long long)j * row + p * a.ld, buf[j][p]);
    }
  // The failure mask rides the same ring (one byte per trace and tick).
  const unsigned char* fp = FAULTS ? a.fail + n0 + (TP ? (long long)ky * a.T * a.ld : 0ll) : nullptr;
  unsigned fbuf[FAULTS ? D : 1][(V + 3) / 4];
  if constexpr (FAULTS) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      if (j < T) load_fail<V>(fp + (long long)j * a.ld, fbuf[j]);
  }

  // Summary mode only: the wide event code of the previous tick of each trace. Bit 4 of a wide
  // code is always 0 (a field counts one event per tick), which ptxas cannot know; xor-ing that bit
  // into the next tick's depth changes nothing but makes the next tick's threshold compares wait
  // for this tick. Without it ptxas hoists the compares of all D unrolled ticks to the top of the
  // loop body, runs out of its 7 predicate registers and keeps them as bits of a general register
  // (62 LOP3 + 9 P2R per 8 decisions in SASS).
  unsigned taint[V];
#pragma unroll
  for (int v = 0; v < V; ++v) taint[v] = 0u;

  const long long* pf_d = dp + (long long)D * row;                  // next row to prefetch (summary mode)
  const unsigned char* pf_f = FAULTS ? fp + (long long)D * a.ld : nullptr;
  auto tick = [&](const int tt, const int j, const bool prefetch, const int phase) {
    long long n[V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      unsigned long long acc = (unsigned long long)buf[j][0][v];  // sqs.go:56-64, wrapping
#pragma unroll
      for (int p = 1; p < A; ++p) acc += (unsigned long long)buf[j][p][v];
      if constexpr (MODE == MODE_SUMMARY && FMA && !FAULTS3) {
        unsigned lo = (unsigned)acc;
        asm("lop3.b32 %0, %0, %1, 16, 0x78;" : "+r"(lo) : "r"(taint[v]));  // lo ^ (taint & 16)
        acc = (acc & 0xffffffff00000000ull) | lo;
      }
      n[v] = (long long)acc;
    }
    unsigned fw[(V + 3) / 4];
    if constexpr (FAULTS) {
#pragma unroll
      for (int w = 0; w < (V + 3) / 4; ++w) fw[w] = fbuf[j][w];
    }
    // Full mode refills the ring slot before the step (the loads are as early as they can be);
    // summary mode refills it after the step: there the step still reads the slot's registers (the
    // high word of the depth in the 64-bit compares) after this point, so ptxas would load into
    // fresh registers and copy the whole ring back at the end of the loop body (17 moves per 8
    // decisions in the first build of the failure-aware kernel).
    auto refill = [&]() {
      if (!prefetch) return;
      if constexpr (MODE == MODE_SUMMARY) {
        // Prefetches happen in tick order, so a running pointer per array replaces the 64-bit
        // multiply per load (with one trace per thread that was ~5 instructions per decision).
#pragma unroll
        for (int p = 0; p < A; ++p) load_depth<V, HINT>(pf_d + p * a.ld, buf[j][p]);
        pf_d += row;
        if constexpr (FAULTS) {
          load_fail<V>(pf_f, fbuf[j]);
          pf_f += a.ld;
        }
      } else {
#pragma unroll
        for (int p = 0; p < A; ++p)
          load_depth<V, HINT>(dp + (long long)(tt + D) * row + p * a.ld, buf[j][p]);
        if constexpr (FAULTS) load_fail<V>(fp + (long long)(tt + D) * a.ld, fbuf[j]);
      }
    };
    if constexpr (MODE != MODE_SUMMARY) refill();
    // tt = tick of the call (0-based) for addressing; i = the tick number the step sees (1-based,
    // relative to the current group of D when REBASE)
    const int i = REBASE ? j + 1 : tt + 1;
    int rr[V];
    unsigned ev[V];  // the event byte; in summary mode the wide event code (ksa_wide_event)
    unsigned fv[V];
    constexpr bool WIDE = MODE == MODE_SUMMARY;
#pragma unroll
    for (int v = 0; v < V; ++v) {
      fv[v] = 0u;
      if constexpr (FAULTS) {
        fv[v] = V == 1 ? fw[0] : (fw[v / 4] >> (8 * (v % 4))) & 0xffu;  // a u8 load is zero-extended
        if constexpr (FAULTS3)  // summary mode only; STEP 9 = ordered thresholds
          ev[v] = ksa_step_faults_fma3<STEP == 9>(c, n[v], fv[v], i, one, s[v], sm[v].pkf);
        else if constexpr (STEP == 7)  // ordered thresholds (summary mode only)
          ev[v] = ksa_step_faults_fma2<WIDE, true>(c, n[v], fv[v], i, one, s[v]);
        else if constexpr ((STEP == 5 && WIDE) || STEP == 8)
          ev[v] = ksa_step_faults_fma2<WIDE, false>(c, n[v], fv[v], i, one, s[v]);
        else if constexpr (STEP == 5)
          ev[v] = ksa_step_faults_fma<false>(c, n[v], fv[v], i, one, s[v]);
        else
          ev[v] = ksa_step_faults(c, n[v], fv[v], i, s[v]);
      } else if constexpr (STEP == 6)
        ev[v] = ksa_step_fma<WIDE, true>(c, n[v], i, one, s[v]);
      else if constexpr (STEP >= 2 && !WIDE)
        ev[v] = ksa_step_fma_byte(c, n[v], i, one, s[v]);
      else if constexpr (STEP >= 2)
        ev[v] = ksa_step_fma<true>(c, n[v], i, one, s[v]);
      else if constexpr (STEP == 1)
        ev[v] = ksa_step_ptx(c, n[v], i, s[v]);
      else
        ev[v] = ksa_step(c, n[v], i, s[v]);
      if constexpr (WIDE && !FMA) ev[v] = ksa_wide_event(ev[v]);
      taint[v] = ev[v];
      rr[v] = s[v].r;
    }
    if constexpr (MODE != MODE_SUMMARY) {
      if (want_r) store_replicas<V>(rp + (long long)tt * a.ld, rr);
      if (want_e) store_events<V, FMA>(ep + (long long)tt * a.ld, ev);
    } else {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        ksa_summary_tick<FMA>(sm[v], rr[v], ev[v], one, phase);
        if constexpr (FAULTS && !FAULTS3) ksa_summary_tick_faults(sm[v], ev[v], fv[v]);
      }
      refill();
    }
  };
  // REBASE: after a group of D ticks, "tick D + k of this group" is "tick k of the next", so both
  // stamps move back by D (one IMAD each per D ticks instead of a tick counter per tick). Stamps stay
  // within int32: >= -(T + 2D) >= -2^30 - 32, <= 2D + KSA_COOL_CAP.
  auto rebase = [&]() {
    if constexpr (REBASE) {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        s[v].su = (int)ksa_mad_u32((unsigned)one, (unsigned)(-D), (unsigned)s[v].su);
        s[v].sd = (int)ksa_mad_u32((unsigned)one, (unsigned)(-D), (unsigned)s[v].sd);
      }
    }
  };

  int t = 0;
  int pending = 0;  // ticks accumulated in the 5-bit counters
  int folds = 0;    // folds accumulated in the 10-bit counters
  // Steady state: every tick of the group exists and so does the tick D ahead of it.
#pragma unroll 1
  for (; t + 2 * D <= T; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j)  // with an even ring depth two ticks share their min / max update
      tick(t + j, j, true, D % 2 ? KSA_TICK_SINGLE : (j % 2 ? KSA_TICK_SECOND : KSA_TICK_FIRST));
    rebase();
    if constexpr (MODE == MODE_SUMMARY) {
      pending += D;
      if (pending > KSA_FOLD_TICKS - D) {
        pending = 0;
        if (++folds == KSA_FLUSH_FOLDS) {
          flush_counters();
          folds = 0;
        } else {
          fold_counters();
        }
      }
    }
  }
  fold_counters();  // the drain adds up to 2D-1 ticks
  // Drain: at most 2D-1 ticks left; guard both the tick and its prefetch.
#pragma unroll 1
  for (; t < T; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      if (t + j < T) tick(t + j, j, t + j + D < T, KSA_TICK_SINGLE);
    rebase();
    fold_counters();
  }
  if constexpr (REBASE) {  // back to ticks of the call: every group took D off, t in total
#pragma unroll
    for (int v = 0; v < V; ++v) {
      s[v].su += t;
      s[v].sd += t;
    }
  }

  if constexpr (MODE == MODE_SUMMARY) {
    flush_counters();
#pragma unroll
    for (int v = 0; v < V; ++v) {
      if (sm[v].bad) {  // a replica count left +-KSA_SUM_SAFE: redo the sum exactly (ksa_summary_acc)
        ksa_trace_regs s0;  // the state this call started from is still in global memory
        s0.r = a.st_r[si + v];
        s0.su = ksa_local_stamp(a.pk.cu, c.cu, tick0, a.st_ju[si + v]);
        s0.sd = ksa_local_stamp(a.pk.cd, c.cd, tick0, a.st_jd[si + v]);
        sm[v].sum = a.summary[si + v].sum_replicas +
                    ksa_exact_sum(a.pk, dp + v, row, a.ld, A, FAULTS ? fp + v : nullptr, a.ld, T, s0);
      }
    }
  }
#pragma unroll
  for (int v = 0; v < V; ++v) {
    a.st_r[si + v] = s[v].r;
    a.st_ju[si + v] = ksa_global_stamp(s[v].su, c.cu, tick0, a.st_ju[si + v]);
    a.st_jd[si + v] = ksa_global_stamp(s[v].sd, c.cd, tick0, a.st_jd[si + v]);
  }
  if constexpr (MODE == MODE_SUMMARY) {
#pragma unroll
    for (int v = 0; v < V; ++v) {
      ksa_summary* q = a.summary + si + v;
      q->final_replicas = s[v].r;
      q->min_replicas = sm[v].rmin;
      q->max_replicas = sm[v].rmax;
      q->sum_replicas = sm[v].sum;
      q->hash = ((unsigned long long)sm[v].hb << 32) | sm[v].ha;
    }
  }
}

#ifdef KSA_TUNING_VARIANTS
// ---- summary replay with the depth rows staged in shared memory by the bulk-copy engine ---------
//
// The register-ring kernel above spends ~3 instructions per decision on load addresses and can
// only keep D * 8 bytes per thread in flight. Here one thread per block issues whole rows
// (BLOCK traces x 8 B = 1 KB, contiguous in the tick-major layout) as cp.async.bulk copies into a
// ring of S stages of K ticks each; completion is signalled on one mbarrier per stage
// (complete_tx), consumption on a second one (one arrive per warp), so no __syncthreads in the
// loop. Every thread then reads its own 8 bytes per tick from shared memory at a compile-time
// offset (conflict-free LDS.64) and runs the same step and summary accumulation as above.
// Block = BLOCK adjacent traces, one per thread; needs n_count to be a multiple of BLOCK (the
// ragged tail goes to the scalar kernel) and the usual alignment (ld % 16 == 0).
__device__ __forceinline__ unsigned ksa_smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void ksa_mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ksa_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void ksa_mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ksa_smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void ksa_mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ksa_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ksa_mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "KSA_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra KSA_DONE;\n\t"
      "bra KSA_WAIT;\n\t"
      "KSA_DONE:\n\t"
      "}" ::"r"(ksa_smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void ksa_bulk_g2s(void* dst_smem, const void* src, unsigned bytes,
                                             unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   ksa_smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(ksa_smem_u32(bar))
               : "memory");
}

// Launched with BLOCK + 32 threads: warps 0..BLOCK/32-1 replay, the last warp only feeds the ring.
template <int A, int K, int S, int BLOCK, int MINB, bool ORDERED>
__global__ void __launch_bounds__(BLOCK + 32, MINB) ksa_summary_bulk_kernel(const __grid_constant__ ReplayArgs a) {
  static_assert(K <= KSA_FOLD_TICKS, "a stage must fit one fold period");
  __shared__ __align__(128) long long ring[S][K][A][BLOCK];
  __shared__ __align__(8) unsigned long long full_bar[S], empty_bar[S];
  const int tid = threadIdx.x;
  const long long nb = a.n_begin + (long long)blockIdx.x * BLOCK;  // first trace of the block
  const long long n = nb + tid;
  const ksa_step_consts c(a.pk);
  const int T = a.T;
  const int one = a.one;

  const int chunks = (T + K - 1) / K;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < S; ++i) {
      ksa_mbar_init(&full_bar[i], 1);
      ksa_mbar_init(&empty_bar[i], BLOCK / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const long long row = (long long)A * a.ld;
  const long long* const src0 = a.depth + nb;
  auto issue = [&](int ch) {  // producer lane: rows of chunk ch into stage ch % S
    const int st = ch % S, t0 = ch * K;
    const int kk = min(K, T - t0);
    ksa_mbar_expect_tx(&full_bar[st], (unsigned)(kk * A * BLOCK * 8));
    for (int k = 0; k < kk; ++k)
#pragma unroll
      for (int p = 0; p < A; ++p)
        ksa_bulk_g2s(&ring[st][k][p][0], src0 + (long long)(t0 + k) * row + p * a.ld, BLOCK * 8, &full_bar[st]);
  };
  if (tid >= BLOCK) {  // producer warp: one lane keeps the ring full, then the warp is done
    if (tid == BLOCK) {
      for (int ch = 0; ch < chunks; ++ch) {
        if (ch >= S) ksa_mbar_wait(&empty_bar[ch % S], (unsigned)((ch / S - 1) & 1));
        issue(ch);
      }
    }
    return;
  }

  ksa_trace_regs s;
  s.r = a.st_r[n];
  s.su = ksa_local_stamp(a.pk.cu, c.cu, a.tick0, a.st_ju[n]);
  s.sd = ksa_local_stamp(a.pk.cd, c.cd, a.tick0, a.st_jd[n]);
  ksa_summary* const q = a.summary + n;
  ksa_summary_acc m;
  ksa_summary_acc_init(m, q->min_replicas, q->max_replicas, q->sum_replicas, q->hash);

  unsigned taint = 0u;  // see ksa_replay_kernel
  auto tick = [&](const int st, const int k, const int i, const int phase) {
    unsigned long long acc = (unsigned long long)ring[st][k][0][tid];  // sqs.go:56-64, wrapping
#pragma unroll
    for (int p = 1; p < A; ++p) acc += (unsigned long long)ring[st][k][p][tid];
    unsigned lo = (unsigned)acc;
    asm("lop3.b32 %0, %0, %1, 16, 0x78;" : "+r"(lo) : "r"(taint));  // lo ^ (taint & 16) == lo
    acc = (acc & 0xffffffff00000000ull) | lo;
    const unsigned w = ksa_step_fma<true, ORDERED>(c, (long long)acc, i, one, s);
    taint = w;
    ksa_summary_tick<true>(m, s.r, w, one, phase);
  };
  int pending = 0, folds = 0;
  const int full_chunks = T / K;
  int st = 0;
  unsigned parity = 0u;
#pragma unroll 1
  for (int ch = 0; ch < full_chunks; ++ch) {  // steady state: every tick of the stage exists
    ksa_mbar_wait(&full_bar[st], parity);
#pragma unroll
    for (int k = 0; k < K; ++k)
      tick(st, k, ch * K + k + 1, K % 2 ? KSA_TICK_SINGLE : (k % 2 ? KSA_TICK_SECOND : KSA_TICK_FIRST));
    __syncwarp();
    if ((tid & 31) == 0) ksa_mbar_arrive(&empty_bar[st]);
    if (++st == S) {
      st = 0;
      parity ^= 1u;
    }
    pending += K;
    if (pending > KSA_FOLD_TICKS - K) {
      pending = 0;
      if (++folds == KSA_FLUSH_FOLDS) {
        ksa_summary_flush(m, q);
        folds = 0;
      } else {
        ksa_summary_fold(m);
      }
    }
  }
  if (full_chunks < chunks) {  // last, partial stage
    ksa_summary_fold(m);
    ksa_mbar_wait(&full_bar[st], parity);
    for (int k = 0; k < T - full_chunks * K; ++k) tick(st, k, full_chunks * K + k + 1, KSA_TICK_SINGLE);
  }

  ksa_summary_flush(m, q);
  if (m.bad) {  // see ksa_replay_kernel
    ksa_trace_regs s0;
    s0.r = a.st_r[n];
    s0.su = ksa_local_stamp(a.pk.cu, c.cu, a.tick0, a.st_ju[n]);
    s0.sd = ksa_local_stamp(a.pk.cd, c.cd, a.tick0, a.st_jd[n]);
    m.sum = q->sum_replicas + ksa_exact_sum(a.pk, a.depth + n, row, a.ld, A, nullptr, 0, T, s0);
  }
  a.st_r[n] = s.r;
  a.st_ju[n] = ksa_global_stamp(s.su, c.cu, a.tick0, a.st_ju[n]);
  a.st_jd[n] = ksa_global_stamp(s.sd, c.cd, a.tick0, a.st_jd[n]);
  q->final_replicas = s.r;
  q->min_replicas = m.rmin;
  q->max_replicas = m.rmax;
  q->sum_replicas = m.sum;
  q->hash = ((unsigned long long)m.hb << 32) | m.ha;
}

#endif  // KSA_TUNING_VARIANTS

typedef void (*replay_kernel_fn)(const ReplayArgs);

// A variant = one tuning point. The register ring costs 2*D*V*A registers, so the multi-attribute
// kernels use a shallower ring / fewer traces per thread than the A = 1 kernel of the same id.
struct Variant {
  int id;
  int step;            // STEP template argument (6 = only valid for ordered policies)
  int tile;            // > 0: the kernel needs n_count to be a multiple of `tile` traces (bulk kernels)
  int V[3], block[3];  // per A-1
  const char* name;
  replay_kernel_fn fn[3];  // indexed by A-1
};

#define KSA_K(A, V, D, HINT, MODE, STEP, BLOCK, MINB) ksa_replay_kernel<A, V, D, HINT, MODE, STEP, BLOCK, MINB>
// A=1 tuning point (V, D, HINT, STEP, BLOCK, MINB); A=2 and A=3 points (V, D, BLOCK, MINB) share HINT/STEP.
#define KSA_VARIANT(ID, MODE, V, D, HINT, STEP, BLOCK, MINB, V2, D2, B2, M2, V3, D3, B3, M3)   \
  {ID, STEP, 0, {V, V2, V3}, {BLOCK, B2, B3}, "V" #V "_D" #D "_H" #HINT "_S" #STEP "_B" #BLOCK "x" #MINB, \
   {KSA_K(1, V, D, HINT, MODE, STEP, BLOCK, MINB), KSA_K(2, V2, D2, HINT == HINT_NA_EF && V2 < 4 ? HINT_NA : HINT, MODE, STEP, B2, M2),                  \
    KSA_K(3, V3, D3, HINT == HINT_NA_EF && V3 < 4 ? HINT_NA : HINT, MODE, STEP, B3, M3)}}

// Bulk-copy summary kernels: (K ticks per stage, S stages) per A; STEP 2 = any policy, 6 = ordered.
#define KSA_BK(A, K, S, BLOCK, MINB, STEP) ksa_summary_bulk_kernel<A, K, S, BLOCK, MINB, STEP == 6>
#define KSA_BULK_VARIANT(ID, STEP, BLOCK, K1, S1, M1, K2, S2, M2, K3, S3, M3)                     \
  {ID, STEP, BLOCK, {1, 1, 1}, {BLOCK + 32, BLOCK + 32, BLOCK + 32}, "BULK_K" #K1 "_S" #S1 "_S" #STEP "_B" #BLOCK "x" #M1, \
   {KSA_BK(1, K1, S1, BLOCK, M1, STEP), KSA_BK(2, K2, S2, BLOCK, M2, STEP), KSA_BK(3, K3, S3, BLOCK, M3, STEP)}}

// Variant 0 is the scalar kernel (any alignment, any N); it also handles the ragged tail of the
// vector variants. The defaults were chosen from tools/sweep_variants.py measurements.
// The product build carries only the kernels replay_common can select by itself: the scalar
// kernels, the defaults per trace count / policy kind, and the optional-output kernels. Every other
// tuning point that was measured (profiles/r01_sweep_variants_*.txt) is compiled only with
// -DKSA_TUNING_VARIANTS (python __graft_entry__.py --tuning), which tools/sweep_variants.py and
// the all-variants parity runs use; ids are stable across the two builds.
const Variant kFullVariants[] = {
    KSA_VARIANT(0, MODE_FULL_OPT, 1, 8, HINT_NA, 0, 128, 4, 1, 4, 128, 4, 1, 4, 128, 4),
    KSA_VARIANT(16, MODE_FULL_OPT, 8, 2, HINT_NA, 2, 256, 2, 8, 1, 128, 4, 8, 1, 128, 3),
    KSA_VARIANT(32, MODE_FULL, 8, 2, HINT_NA, 2, 256, 2, 8, 1, 128, 4, 8, 1, 128, 3),
    // fewer traces per thread: for trace counts that cannot fill the GPU at V = 8 (see replay_common)
    KSA_VARIANT(42, MODE_FULL, 1, 8, HINT_NA, 2, 128, 8, 1, 4, 128, 8, 1, 3, 128, 8),
    KSA_VARIANT(43, MODE_FULL, 2, 4, HINT_NA, 2, 128, 8, 2, 3, 128, 8, 2, 2, 128, 8),
    KSA_VARIANT(46, MODE_FULL_OPT, 1, 8, HINT_NA, 2, 128, 8, 1, 4, 128, 8, 1, 3, 128, 8),
#ifdef KSA_TUNING_VARIANTS
    KSA_VARIANT(1, MODE_FULL, 4, 4, HINT_NA, 0, 128, 7, 4, 2, 128, 6, 2, 4, 128, 6),
    KSA_VARIANT(14, MODE_FULL, 4, 3, HINT_NA, 1, 128, 8, 4, 2, 128, 6, 2, 3, 128, 7),
    KSA_VARIANT(19, MODE_FULL, 4, 3, HINT_NA, 1, 256, 4, 4, 2, 256, 3, 2, 3, 256, 3),
    KSA_VARIANT(20, MODE_FULL, 2, 4, HINT_NA, 1, 128, 8, 2, 3, 128, 8, 2, 2, 128, 8),
    KSA_VARIANT(21, MODE_FULL, 2, 6, HINT_NA, 1, 256, 4, 2, 3, 256, 4, 2, 2, 256, 4),
    KSA_VARIANT(23, MODE_FULL, 8, 2, HINT_NA, 1, 128, 4, 4, 3, 128, 5, 4, 2, 128, 5),
    KSA_VARIANT(27, MODE_FULL, 1, 8, HINT_NA, 1, 128, 8, 1, 4, 128, 8, 1, 3, 128, 8),
    KSA_VARIANT(30, MODE_FULL, 8, 2, HINT_NA, 2, 128, 4, 4, 3, 128, 5, 4, 2, 128, 5),
    KSA_VARIANT(31, MODE_FULL, 8, 3, HINT_NA, 2, 128, 3, 4, 4, 128, 4, 4, 3, 128, 4),
    KSA_VARIANT(34, MODE_FULL, 8, 1, HINT_NA, 2, 128, 5, 8, 1, 128, 5, 4, 1, 128, 6),
    KSA_VARIANT(36, MODE_FULL, 16, 2, HINT_NA, 2, 64, 4, 8, 2, 64, 6, 8, 1, 64, 8),
    KSA_VARIANT(38, MODE_FULL, 8, 2, HINT_NA_EF, 2, 128, 4, 4, 3, 128, 5, 4, 2, 128, 5),
    KSA_VARIANT(41, MODE_FULL, 8, 2, HINT_NA, 0, 128, 4, 4, 3, 128, 5, 4, 2, 128, 5),
    KSA_VARIANT(44, MODE_FULL, 4, 3, HINT_NA, 2, 128, 8, 4, 2, 128, 6, 2, 3, 128, 7),
    KSA_VARIANT(45, MODE_FULL, 1, 8, HINT_NA, 2, 64, 16, 1, 4, 64, 16, 1, 3, 64, 16),
#endif
};
const Variant kSummaryVariants[] = {
    KSA_VARIANT(0, MODE_SUMMARY, 1, 8, HINT_NA, 0, 128, 4, 1, 4, 128, 4, 1, 4, 128, 4),
    // ring of 12 ticks: 1.000 (16) and 1.005 (24) of the measured copy peak at A = 1 against 0.979 / 0.987
    // for the ring of 8 (20, 21) once the loop is at 32 instructions per decision
    // (profiles/r02_variants.txt); A = 2, 3 keep the shapes of 20 / 21
    KSA_VARIANT(16, MODE_SUMMARY, 1, 12, HINT_NA, 2, 128, 8, 1, 3, 128, 10, 1, 2, 128, 10),
    // STEP 6: for policies with scale_up_messages > scale_down_messages only (ksa_step_fma ORDERED);
    // replay_common substitutes the general default when the policy is not ordered.
    KSA_VARIANT(24, MODE_SUMMARY, 1, 12, HINT_NA, 6, 128, 8, 1, 3, 128, 10, 1, 2, 128, 10),
#ifdef KSA_TUNING_VARIANTS
    KSA_VARIANT(20, MODE_SUMMARY, 1, 8, HINT_NA, 2, 128, 8, 1, 3, 128, 10, 1, 2, 128, 10),
    KSA_VARIANT(21, MODE_SUMMARY, 1, 8, HINT_NA, 6, 128, 8, 1, 3, 128, 10, 1, 2, 128, 10),
    KSA_VARIANT(5, MODE_SUMMARY, 2, 8, HINT_NA, 1, 128, 4, 2, 4, 128, 4, 2, 3, 128, 4),
    KSA_VARIANT(7, MODE_SUMMARY, 1, 8, HINT_NA, 2, 128, 8, 1, 4, 128, 8, 1, 3, 128, 8),
    KSA_VARIANT(8, MODE_SUMMARY, 2, 8, HINT_NA, 2, 128, 4, 2, 4, 128, 4, 2, 3, 128, 4),
    KSA_VARIANT(9, MODE_SUMMARY, 4, 4, HINT_NA, 2, 128, 4, 4, 2, 128, 4, 4, 2, 128, 3),
    KSA_VARIANT(10, MODE_SUMMARY, 8, 2, HINT_NA, 2, 128, 3, 4, 3, 128, 3, 4, 2, 128, 3),
    KSA_VARIANT(12, MODE_SUMMARY, 2, 4, HINT_NA, 2, 128, 8, 2, 3, 128, 8, 2, 2, 128, 8),
    KSA_VARIANT(13, MODE_SUMMARY, 1, 4, HINT_NA, 2, 128, 12, 1, 2, 128, 12, 1, 2, 128, 10),
    KSA_VARIANT(14, MODE_SUMMARY, 1, 6, HINT_NA, 2, 128, 10, 1, 3, 128, 10, 1, 2, 128, 10),
    KSA_VARIANT(15, MODE_SUMMARY, 1, 8, HINT_NA, 2, 256, 4, 1, 4, 256, 4, 1, 3, 256, 4),
    KSA_VARIANT(31, MODE_SUMMARY, 1, 12, HINT_NA, 2, 128, 8, 1, 6, 128, 8, 1, 4, 128, 8),
    KSA_VARIANT(17, MODE_SUMMARY, 1, 16, HINT_NA, 2, 128, 6, 1, 8, 128, 6, 1, 5, 128, 6),
    KSA_VARIANT(18, MODE_SUMMARY, 1, 8, HINT_NA, 2, 128, 10, 1, 4, 128, 10, 1, 3, 128, 10),
    KSA_VARIANT(19, MODE_SUMMARY, 1, 8, HINT_NA, 2, 64, 16, 1, 4, 64, 16, 1, 3, 64, 16),
    KSA_VARIANT(22, MODE_SUMMARY, 2, 4, HINT_NA, 6, 128, 8, 2, 3, 128, 8, 2, 2, 128, 8),
    KSA_VARIANT(23, MODE_SUMMARY, 1, 8, HINT_NA, 6, 128, 10, 1, 4, 128, 10, 1, 3, 128, 10),
    KSA_VARIANT(32, MODE_SUMMARY, 1, 12, HINT_NA, 6, 128, 8, 1, 6, 128, 8, 1, 4, 128, 8),
    KSA_VARIANT(25, MODE_SUMMARY, 1, 8, HINT_NA, 6, 256, 4, 1, 4, 256, 4, 1, 3, 256, 4),
    // 30..: rows staged in shared memory by cp.async.bulk (ksa_summary_bulk_kernel). Measured
    // 0.87 (33) / 0.86 (38) of the copy bandwidth against 0.92 for the register ring (21), so they
    // are tuning points, not the default; K = 4 stages reach 0.77 (profiles/r01_summary_bulk.txt).
    KSA_BULK_VARIANT(30, 2, 128, 8, 3, 8, 4, 3, 8, 2, 4, 8),
    KSA_BULK_VARIANT(33, 6, 128, 8, 3, 8, 4, 3, 8, 2, 4, 8),
    KSA_BULK_VARIANT(38, 6, 128, 8, 4, 6, 4, 4, 6, 2, 4, 6),
#endif
};
// The *_faults entry points (STEP 4 = the C++ step: the scalar kernels and the optional-output one).
const Variant kFullFaultsVariants[] = {
    KSA_VARIANT(0, MODE_FULL_OPT, 1, 8, HINT_NA, 4, 128, 4, 1, 4, 128, 4, 1, 4, 128, 4),
    KSA_VARIANT(1, MODE_FULL_OPT, 8, 2, HINT_NA, 4, 128, 4, 4, 3, 128, 4, 4, 2, 128, 4),
    // STEP 8: the second spelling of the failure-aware step (ksa_step_faults_fma2), event byte out
    KSA_VARIANT(8, MODE_FULL, 4, 4, HINT_NA, 8, 128, 5, 4, 2, 128, 5, 2, 3, 128, 6),
    // one trace per thread below 262144 traces: 0.83 / 0.82 / 0.77 of the copy peak at 98304 / 131072 / 196608
    // against 0.49 / 0.62 / 0.74 for four per thread (profiles/r02_small_n_modes.txt)
    KSA_VARIANT(9, MODE_FULL, 1, 8, HINT_NA, 8, 128, 8, 1, 4, 128, 8, 1, 3, 128, 8),
#ifdef KSA_TUNING_VARIANTS
    KSA_VARIANT(2, MODE_FULL, 8, 2, HINT_NA, 5, 128, 4, 4, 3, 128, 4, 4, 2, 128, 4),
    KSA_VARIANT(3, MODE_FULL, 8, 2, HINT_NA, 5, 256, 2, 4, 3, 256, 2, 4, 2, 256, 2),
    KSA_VARIANT(4, MODE_FULL, 4, 3, HINT_NA, 5, 128, 6, 4, 2, 128, 5, 2, 3, 128, 6),
    KSA_VARIANT(5, MODE_FULL, 4, 3, HINT_NA, 8, 128, 6, 4, 2, 128, 5, 2, 3, 128, 6),
    KSA_VARIANT(6, MODE_FULL, 8, 2, HINT_NA, 8, 128, 4, 4, 3, 128, 4, 4, 2, 128, 4),
#endif
};
const Variant kSummaryFaultsVariants[] = {
    KSA_VARIANT(0, MODE_SUMMARY, 1, 8, HINT_NA, 4, 128, 4, 1, 4, 128, 4, 1, 4, 128, 4),
    // STEP 9 / 10: ksa_step_faults_fma3 (counts the failures itself); 9 = ordered thresholds only,
    // replay_common substitutes 10 otherwise
    // ring of 12 ticks, 7 blocks of 128 per SM (73 registers): 6.45 ms on 2^20 x 4000 with failures =
    // 0.906 of the measured copy peak; the ring of 8 at 8 blocks per SM (ids 10 / 11): 6.66 ms, 0.878
    // (profiles/r02_faults_tune.txt). A = 2, 3 keep the shapes of 10 / 11.
    KSA_VARIANT(28, MODE_SUMMARY, 1, 12, HINT_NA, 10, 128, 7, 1, 3, 128, 8, 1, 2, 128, 8),
    KSA_VARIANT(24, MODE_SUMMARY, 1, 12, HINT_NA, 9, 128, 7, 1, 3, 128, 8, 1, 2, 128, 8),
#ifdef KSA_TUNING_VARIANTS
    KSA_VARIANT(10, MODE_SUMMARY, 1, 8, HINT_NA, 10, 128, 8, 1, 3, 128, 8, 1, 2, 128, 8),
    KSA_VARIANT(11, MODE_SUMMARY, 1, 8, HINT_NA, 9, 128, 8, 1, 3, 128, 8, 1, 2, 128, 8),
    KSA_VARIANT(1, MODE_SUMMARY, 2, 4, HINT_NA, 4, 128, 6, 2, 3, 128, 6, 2, 2, 128, 6),
    KSA_VARIANT(2, MODE_SUMMARY, 2, 4, HINT_NA, 5, 128, 7, 2, 3, 128, 6, 2, 2, 128, 6),
    KSA_VARIANT(3, MODE_SUMMARY, 2, 4, HINT_NA, 5, 128, 6, 2, 3, 128, 6, 2, 2, 128, 6),
    KSA_VARIANT(4, MODE_SUMMARY, 1, 8, HINT_NA, 5, 128, 8, 1, 3, 128, 8, 1, 2, 128, 8),
    KSA_VARIANT(5, MODE_SUMMARY, 1, 8, HINT_NA, 5, 128, 6, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(6, MODE_SUMMARY, 1, 6, HINT_NA, 5, 128, 10, 1, 3, 128, 10, 1, 2, 128, 10),
    // STEP 7: ksa_step_faults_fma2, ordered thresholds only (4 is its twin for any policy)
    KSA_VARIANT(7, MODE_SUMMARY, 1, 8, HINT_NA, 7, 128, 8, 1, 3, 128, 8, 1, 2, 128, 8),
    KSA_VARIANT(8, MODE_SUMMARY, 2, 4, HINT_NA, 7, 128, 6, 2, 3, 128, 6, 2, 2, 128, 6),
    KSA_VARIANT(9, MODE_SUMMARY, 1, 8, HINT_NA, 7, 128, 6, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(12, MODE_SUMMARY, 1, 6, HINT_NA, 9, 128, 10, 1, 3, 128, 10, 1, 2, 128, 10),
    KSA_VARIANT(13, MODE_SUMMARY, 1, 12, HINT_NA, 9, 128, 6, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(14, MODE_SUMMARY, 2, 4, HINT_NA, 9, 128, 6, 2, 3, 128, 6, 2, 2, 128, 6),
    KSA_VARIANT(15, MODE_SUMMARY, 1, 8, HINT_NA, 9, 128, 6, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(16, MODE_SUMMARY, 1, 8, HINT_NA, 9, 256, 4, 1, 4, 256, 4, 1, 3, 256, 4),
    KSA_VARIANT(17, MODE_SUMMARY, 1, 8, HINT_NA, 9, 128, 10, 1, 4, 128, 10, 1, 3, 128, 10),
    KSA_VARIANT(22, MODE_SUMMARY, 1, 16, HINT_NA, 9, 128, 5, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(23, MODE_SUMMARY, 1, 24, HINT_NA, 9, 128, 4, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(29, MODE_SUMMARY, 1, 12, HINT_NA, 9, 128, 7, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(30, MODE_SUMMARY, 1, 10, HINT_NA, 9, 128, 7, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(31, MODE_SUMMARY, 1, 14, HINT_NA, 9, 128, 6, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(25, MODE_SUMMARY, 1, 12, HINT_NA, 9, 256, 3, 1, 4, 256, 3, 1, 3, 256, 3),
    KSA_VARIANT(26, MODE_SUMMARY, 1, 12, HINT_NA, 9, 128, 5, 1, 4, 128, 6, 1, 3, 128, 6),
    KSA_VARIANT(27, MODE_SUMMARY, 1, 16, HINT_NA, 9, 128, 4, 1, 4, 128, 6, 1, 3, 128, 6),
#endif
};
// The summary defaults (16 / 24, with failures 28 / 24) as chunk-of-the-time-axis kernels for
// ksa_timepar.cu: [failure mask * 2 + ordered thresholds][A - 1]. One trace per thread, scalar
// loads: any layout.
#define KSA_KTP(A, D, STEP, MINB) ksa_replay_kernel<A, 1, D, HINT_NA, MODE_SUMMARY, STEP, 128, MINB, KSA_CHUNKS>
#define KSA_KTL(A, D, STEP, MINB) ksa_replay_kernel<A, 1, D, HINT_NA, MODE_SUMMARY, STEP, 128, MINB, KSA_CHUNKS_MARKED>
const replay_kernel_fn kSummaryChunkKernels[4][3] = {
    {KSA_KTP(1, 12, 2, 8), KSA_KTP(2, 3, 2, 10), KSA_KTP(3, 2, 2, 10)},
    {KSA_KTP(1, 12, 6, 8), KSA_KTP(2, 3, 6, 10), KSA_KTP(3, 2, 6, 10)},
    {KSA_KTP(1, 12, 10, 7), KSA_KTP(2, 3, 10, 8), KSA_KTP(3, 2, 10, 8)},
    {KSA_KTP(1, 12, 9, 7), KSA_KTP(2, 3, 9, 8), KSA_KTP(3, 2, 9, 8)},
};
// ... and for marked chunks only (the chunks ksa_timepar.cu replays a second time)
const replay_kernel_fn kSummaryMarkedChunkKernels[4][3] = {
    {KSA_KTL(1, 12, 2, 8), KSA_KTL(2, 3, 2, 10), KSA_KTL(3, 2, 2, 10)},
    {KSA_KTL(1, 12, 6, 8), KSA_KTL(2, 3, 6, 10), KSA_KTL(3, 2, 6, 10)},
    {KSA_KTL(1, 12, 10, 7), KSA_KTL(2, 3, 10, 8), KSA_KTL(3, 2, 10, 8)},
    {KSA_KTL(1, 12, 9, 7), KSA_KTL(2, 3, 9, 8), KSA_KTL(3, 2, 9, 8)},
};
// The one-trace-per-thread defaults of the four entry points as gated kernels (only the traces flagged in
// ReplayArgs::gate): what follows the chunked passes. Summary: the table above with KSA_GATED; full mode:
// variants 42 and, with a failure mask, 9. [A - 1]
#define KSA_KG(A, D, MODE, STEP, MINB) ksa_replay_kernel<A, 1, D, HINT_NA, MODE, STEP, 128, MINB, KSA_GATED>
const replay_kernel_fn kSummaryGatedKernels[4][3] = {
    {KSA_KG(1, 12, MODE_SUMMARY, 2, 8), KSA_KG(2, 3, MODE_SUMMARY, 2, 10), KSA_KG(3, 2, MODE_SUMMARY, 2, 10)},
    {KSA_KG(1, 12, MODE_SUMMARY, 6, 8), KSA_KG(2, 3, MODE_SUMMARY, 6, 10), KSA_KG(3, 2, MODE_SUMMARY, 6, 10)},
    {KSA_KG(1, 12, MODE_SUMMARY, 10, 7), KSA_KG(2, 3, MODE_SUMMARY, 10, 8), KSA_KG(3, 2, MODE_SUMMARY, 10, 8)},
    {KSA_KG(1, 12, MODE_SUMMARY, 9, 7), KSA_KG(2, 3, MODE_SUMMARY, 9, 8), KSA_KG(3, 2, MODE_SUMMARY, 9, 8)},
};
const replay_kernel_fn kFullGatedKernels[2][3] = {
    {KSA_KG(1, 8, MODE_FULL, 2, 8), KSA_KG(2, 4, MODE_FULL, 2, 8), KSA_KG(3, 3, MODE_FULL, 2, 8)},
    {KSA_KG(1, 8, MODE_FULL, 8, 8), KSA_KG(2, 4, MODE_FULL, 8, 8), KSA_KG(3, 3, MODE_FULL, 8, 8)},
};
constexpr int kDefaultFull = 32;      // profiles/r01_sweep_variants_a1_T2000.txt: 1.007 of the measured copy peak
constexpr int kDefaultFullV1 = 42;   // fewer than 196608 traces: one trace per thread
constexpr int kDefaultFullV2 = 43;   // fewer than 327680 traces: two traces per thread
constexpr int kDefaultFullOpt = 16;  // used when one of the two outputs is NULL
constexpr int kDefaultFullOptV1 = 46;  // the same below 196608 traces
constexpr int kDefaultSummary = 16;  // one trace per thread: no interleaved state machines, no predicate spills
constexpr int kDefaultSummaryOrdered = 24;  // the same with the ordered-thresholds step
constexpr int kDefaultFullFaults = 8, kDefaultFullFaultsV1 = 9, kDefaultFullFaultsOpt = 1;
constexpr int kDefaultSummaryFaults = 28, kDefaultSummaryFaultsOrdered = 24;

template <size_t K>
const Variant* find_variant(const Variant (&table)[K], int id) {
  for (size_t i = 0; i < K; ++i)
    if (table[i].id == id) return &table[i];
  return nullptr;
}

int launch(const Variant& var, int A, ReplayArgs args, cudaStream_t stream) {
  if (args.n_count <= 0) return KSA_OK;
  const long long threads = (args.n_count + var.V[A - 1] - 1) / var.V[A - 1];
  const long long blocks = var.tile > 0 ? args.n_count / var.tile : (threads + var.block[A - 1] - 1) / var.block[A - 1];
  if (blocks > 0x7fffffffll) return ksa_fail(KSA_ELIMIT, "grid too large (%lld blocks)", blocks);
  var.fn[A - 1]<<<dim3((unsigned)blocks), dim3((unsigned)var.block[A - 1]), 0, stream>>>(args);
  KSA_LAUNCH_CHECK(var.name);
  return KSA_OK;
}

int replay_common(int mode, bool faults, const ksa_policy* policy, const int64_t* depth_dev,
                  const uint8_t* fail_dev, int64_t T, int32_t A, int64_t N, int64_t ld, int64_t tick0,
                  const ksa_state* state, int32_t* replicas_dev, uint8_t* events_dev,
                  ksa_summary* summary_dev, void* stream) {
  if (!policy || !depth_dev || !state) return ksa_fail(KSA_EINVAL, "NULL policy, depth or state");
  if (faults && !fail_dev) return ksa_fail(KSA_EINVAL, "NULL fail_dev");
  if (!state->replicas_dev || !state->last_up_tick_dev || !state->last_down_tick_dev)
    return ksa_fail(KSA_EINVAL, "NULL pointer inside ksa_state");
  if (mode == MODE_SUMMARY && !summary_dev) return ksa_fail(KSA_EINVAL, "NULL summary_dev");
  if (T < 1 || T > KSA_MAX_CHUNK_TICKS)
    return ksa_fail(KSA_EINVAL, "T=%lld outside [1, 2^30 - 64]; chunk longer traces", (long long)T);
  if (N < 1 || N > 0x7fffffffll) return ksa_fail(KSA_EINVAL, "N=%lld outside [1, 2^31-1]", (long long)N);
  if (ld < N) return ksa_fail(KSA_EINVAL, "ld=%lld < N=%lld", (long long)ld, (long long)N);
  if (tick0 < 0) return ksa_fail(KSA_EINVAL, "tick0=%lld < 0", (long long)tick0);
  ReplayArgs args;
  int rc = ksa_policy_to_k(policy, A, &args.pk);
  if (rc != KSA_OK) return rc;
  args.depth = (const long long*)depth_dev;
  args.fail = faults ? fail_dev : nullptr;
  args.ld = ld;
  args.tick0 = tick0;
  args.T = (int)T;
  args.one = 1;
  args.st_r = state->replicas_dev;
  args.st_ju = (long long*)state->last_up_tick_dev;
  args.st_jd = (long long*)state->last_down_tick_dev;
  args.replicas = replicas_dev;
  args.events = events_dev;
  args.summary = summary_dev;
  args.tp_total = 0;
  args.tp_stride = 0;
  args.gate = nullptr;
  args.gate_value = 0;

  const int want = mode == MODE_FULL ? ksa_full_variant() : ksa_summary_variant();
  const bool ordered = policy->scale_up_messages > policy->scale_down_messages;
  // Trace counts far below what fills the GPU: chunks of the time axis in parallel (ksa_timepar.cu;
  // same results). Only the library default takes this path; an explicit variant gets its kernel.
  if (want < 0 && (mode == MODE_SUMMARY || (replicas_dev && events_dev)) &&
      !(faults && std::getenv("KSA_FAULTS_VARIANT"))) {
    const bool vector_ok = mode == MODE_FULL && (N % 8 == 0) && (ld % 16 == 0) && ksa_aligned(depth_dev, 32) &&
                           ksa_aligned(replicas_dev, 32) && ksa_aligned(events_dev, 16) &&
                           (!faults || ksa_aligned(fail_dev, 16));
    int L = 0, warm = 0, wide = 0;
    const int K = ksa_timepar_plan(T, N, mode == MODE_SUMMARY, faults, vector_ok, &L, &warm, &wide);
    if (K >= 2) {
      cudaStream_t s = (cudaStream_t)stream;
      // what follows the chunked passes: the one-trace-per-thread default of this call (scalar accesses:
      // any layout) over the traces they flagged
      const replay_kernel_fn gated = mode == MODE_FULL ? kFullGatedKernels[faults ? 1 : 0][A - 1]
                                                       : kSummaryGatedKernels[(faults ? 2 : 0) + (ordered ? 1 : 0)][A - 1];
      const ksa_plain_launch plain = [&](const int* unresolved_dev) {
        ReplayArgs g = args;
        g.gate = unresolved_dev;
        g.n_begin = 0;
        g.n_count = N;
        gated<<<dim3((unsigned)((N + 127) / 128)), 128, 0, s>>>(g);
        KSA_LAUNCH_CHECK("ksa_replay_kernel (traces the chunked passes gave up)");
        return KSA_OK;
      };
      rc = mode == MODE_FULL
               ? ksa_timepar_replay_full(args.pk, depth_dev, args.fail, T, A, N, ld, tick0, state, replicas_dev,
                                         events_dev, K, L, warm, wide, plain, s)
               : ksa_timepar_replay_summary(args.pk, ordered, depth_dev, args.fail, T, A, N, ld, tick0, state,
                                            summary_dev, K, L, warm, plain, s);
      if (rc <= 0) return rc;  // done, or an error; 1 = the scratch is busy: plain kernels below
    }
  }
  const int default_summary = ordered ? kDefaultSummaryOrdered : kDefaultSummary;
  // Full mode: 8 traces per thread need ~600k traces to fill 148 SMs; below that the kernel is bound
  // by the per-thread state chain, not by HBM, and fewer traces per thread win
  // (profiles/r01_small_n_variants.txt: at N = 65536, V = 8: 0.30 of the copy peak, V = 1: 0.73).
  const int default_full = N < 196608 ? kDefaultFullV1 : (N < 327680 ? kDefaultFullV2 : kDefaultFull);
  const Variant* var = mode == MODE_FULL
                           ? find_variant(kFullVariants, want < 0 ? default_full : want)
                           : find_variant(kSummaryVariants, want < 0 ? default_summary : want);
  if (var && var->step == 6 && !ordered) var = find_variant(kSummaryVariants, kDefaultSummary);
  // Variants compiled for "both outputs" cannot run with one of them missing.
  if (!faults && mode == MODE_FULL && (!replicas_dev || !events_dev))
    var = find_variant(kFullVariants, N < 196608 ? kDefaultFullOptV1 : kDefaultFullOpt);
  const Variant* scalar = mode == MODE_FULL ? &kFullVariants[0] : &kSummaryVariants[0];
  if (faults) {
    // The tuning knob KSA_FAULTS_VARIANT=<id> (environment) selects another table entry for measurements.
    const bool both = replicas_dev && events_dev;
    int fid = mode == MODE_FULL ? (both ? (N < 262144 ? kDefaultFullFaultsV1 : kDefaultFullFaults) : kDefaultFullFaultsOpt)
                                : (ordered ? kDefaultSummaryFaultsOrdered : kDefaultSummaryFaults);
    const int fid_default = fid;
    if (const char* e = std::getenv("KSA_FAULTS_VARIANT")) fid = std::atoi(e);
    var = mode == MODE_FULL ? find_variant(kFullFaultsVariants, fid) : find_variant(kSummaryFaultsVariants, fid);
    // one knob serves both modes: an id only the other mode's table has means "this mode's default"
    if (!var) var = mode == MODE_FULL ? find_variant(kFullFaultsVariants, fid_default) : find_variant(kSummaryFaultsVariants, fid_default);
    // steps for ordered thresholds only: substitute the twin for any policy
    if (var->step == 9 && !ordered) var = find_variant(kSummaryFaultsVariants, kDefaultSummaryFaults);
    if (var->step == 7 && !ordered) var = find_variant(kSummaryFaultsVariants, 4);
    // variants compiled for "both outputs" cannot run with one of them missing
    if (mode == MODE_FULL && !both && var->id >= 2) var = find_variant(kFullFaultsVariants, kDefaultFullFaultsOpt);
    if (want == 0) var = mode == MODE_FULL ? &kFullFaultsVariants[0] : &kSummaryFaultsVariants[0];
    scalar = mode == MODE_FULL ? &kFullFaultsVariants[0] : &kSummaryFaultsVariants[0];
  }
  if (!var) return ksa_fail(KSA_EINVAL, "unknown kernel variant %d", want);
  // Vector path needs every row start 32-byte aligned in all three arrays.
  const bool aligned = (ld % 16 == 0) && ksa_aligned(depth_dev, 32) && (!faults || ksa_aligned(fail_dev, 16)) &&
                       (!replicas_dev || ksa_aligned(replicas_dev, 32)) &&
                       (!events_dev || ksa_aligned(events_dev, 16));
  if (!aligned) var = scalar;
  cudaStream_t st = (cudaStream_t)stream;
  const int V = var->tile > 0 ? var->tile : var->V[A - 1];  // granularity of the vector launch
  const long long n_vec = (N / V) * V;
  args.n_begin = 0;
  args.n_count = n_vec;
  rc = launch(*var, A, args, st);
  if (rc != KSA_OK) return rc;
  if (n_vec < N) {  // ragged tail: fewer than V traces, scalar kernel
    args.n_begin = n_vec;
    args.n_count = N - n_vec;
    rc = launch(*scalar, A, args, st);
  }
  return rc;
}

}  // namespace

// Pass 1 of the time-parallel summary replay (ksa_timepar.cu): K chunks of L ticks of every trace,
// each replayed like a call of its own by the default summary kernel, with the chunk's carry state
// at st_*[k * N + n] (in: the state the chunk starts from; out: the state it ends with) and its
// summary accumulated into partial[k * N + n].
// marks_dev != NULL: only the chunks with marks_dev[k * N + n] == mark.
int ksa_replay_summary_chunks(const ksa_policy_k& pk, bool ordered, const int64_t* depth_dev, const uint8_t* fail_dev,
                              int L, int64_t T, int32_t A, int64_t N, int64_t ld, int64_t tick0, int K, int32_t* st_r,
                              int64_t* st_ju, int64_t* st_jd, ksa_summary* partial, const int* marks_dev, int mark,
                              cudaStream_t stream) {
  ReplayArgs args;
  args.pk = pk;
  args.depth = (const long long*)depth_dev;
  args.fail = fail_dev;
  args.ld = ld;
  args.tick0 = tick0;
  args.n_begin = 0;
  args.n_count = N;
  args.T = L;
  args.one = 1;
  args.st_r = st_r;
  args.st_ju = (long long*)st_ju;
  args.st_jd = (long long*)st_jd;
  args.replicas = nullptr;
  args.events = nullptr;
  args.summary = partial;
  args.tp_total = (int)T;
  args.tp_stride = N;
  args.gate = marks_dev;
  args.gate_value = mark;
  const int kind = (fail_dev ? 2 : 0) + (ordered ? 1 : 0);
  const long long blocks = (N + 127) / 128;
  (marks_dev ? kSummaryMarkedChunkKernels : kSummaryChunkKernels)[kind][A - 1]<<<dim3((unsigned)blocks, (unsigned)K), 128, 0, stream>>>(args);
  KSA_LAUNCH_CHECK("ksa_replay_kernel (chunks of the time axis)");
  return KSA_OK;
}

extern "C" int ksa_replay_full(const ksa_policy* policy, const int64_t* depth_dev, int64_t T,
                               int32_t A, int64_t N, int64_t ld, int64_t tick0,
                               const ksa_state* state, int32_t* replicas_dev, uint8_t* events_dev,
                               void* stream) {
  return replay_common(MODE_FULL, false, policy, depth_dev, nullptr, T, A, N, ld, tick0, state,
                       replicas_dev, events_dev, nullptr, stream);
}

extern "C" int ksa_replay_summary(const ksa_policy* policy, const int64_t* depth_dev, int64_t T,
                                  int32_t A, int64_t N, int64_t ld, int64_t tick0,
                                  const ksa_state* state, ksa_summary* summary_dev, void* stream) {
  return replay_common(MODE_SUMMARY, false, policy, depth_dev, nullptr, T, A, N, ld, tick0, state,
                       nullptr, nullptr, summary_dev, stream);
}

extern "C" int ksa_replay_full_faults(const ksa_policy* policy, const int64_t* depth_dev,
                                      const uint8_t* fail_dev, int64_t T, int32_t A, int64_t N,
                                      int64_t ld, int64_t tick0, const ksa_state* state,
                                      int32_t* replicas_dev, uint8_t* events_dev, void* stream) {
  return replay_common(MODE_FULL, true, policy, depth_dev, fail_dev, T, A, N, ld, tick0, state,
                       replicas_dev, events_dev, nullptr, stream);
}

extern "C" int ksa_replay_summary_faults(const ksa_policy* policy, const int64_t* depth_dev,
                                         const uint8_t* fail_dev, int64_t T, int32_t A, int64_t N,
                                         int64_t ld, int64_t tick0, const ksa_state* state,
                                         ksa_summary* summary_dev, void* stream) {
  return replay_common(MODE_SUMMARY, true, policy, depth_dev, fail_dev, T, A, N, ld, tick0, state,
                       nullptr, nullptr, summary_dev, stream);
}

// Used by ksa_set_variant to validate ids.
bool ksa_replay_variant_exists(int mode, int id) {
  return mode == MODE_FULL ? find_variant(kFullVariants, id) != nullptr
                           : find_variant(kSummaryVariants, id) != nullptr;
}
// Variant names for tools/ (not part of the C ABI header; looked up with dlsym by the sweep tool).
const char* ksa_sweep_variant_name(int id);  // ksa_sweep.cu
extern "C" const char* ksa_variant_name(int32_t mode, int32_t id) {
  if (mode == 4) return ksa_sweep_variant_name(id);  // KSA_SWEEP_VARIANT ids
  const Variant* v = nullptr;
  switch (mode) {
    case 0: v = find_variant(kFullVariants, id == -1 ? kDefaultFull : id); break;
    case 1: v = find_variant(kSummaryVariants, id == -1 ? kDefaultSummary : id); break;
    case 2: v = find_variant(kFullFaultsVariants, id == -1 ? kDefaultFullFaults : id); break;  // KSA_FAULTS_VARIANT ids
    case 3: v = find_variant(kSummaryFaultsVariants, id == -1 ? kDefaultSummaryFaults : id); break;
    default: break;
  }
  return v ? v->name : nullptr;
}
