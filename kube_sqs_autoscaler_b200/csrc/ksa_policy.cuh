// ksa_policy.cuh — the scaling policy as one branch-free device function, shared by every replay
// kernel, plus the kernel-side policy record and the chunk-local encoding of the carry state.
//
// What it computes per (trace, tick) — reference file:line in /root/reference:
//   threshold tests   n >= scaleUpMessages / n <= scaleDownMessages (independent ifs) main.go:51,65
//   cool-down gates   lastScaleXTime.Add(cool).After(now) => skip the rest of the tick  main.go:52-55,66-69
//   ScaleUp           no-op at r >= int32(Max); r + int32(up), clamp to Max only     scale/scale.go:60-69
//   ScaleDown         no-op at r <= int32(Min); r - int32(down), clamp to Min only   scale/scale.go:87-97
//   stamps            taken whenever Scale* returned nil, also after a no-op          main.go:62,76
// The oracle (oracle/ksa_oracle.c) states the same thing with the reference's control flow; this
// file states it with predicates so that a warp never diverges.
#pragma once

#include <cstdint>
#include <climits>

#include "../../include/ksa_replay.h"

// Kernel-side policy record, 48 bytes. Built on the host by ksa_policy_to_k (ksa_api.cu).
struct ksa_policy_k {
  long long up_msgs;    // scale_up_messages
  long long down_msgs;  // scale_down_messages
  long long cu;         // up cool-down in ticks, max(0, ceil(cool/poll))
  long long cd;         // down cool-down in ticks
  int max32;            // int32(p.Max)          scale/scale.go:62,67,68
  int min32;            // int32(p.Min)          scale/scale.go:89,95,96
  int up32;             // int32(p.ScaleUpPods)  scale/scale.go:66
  int down32;           // int32(p.ScaleDownPods) scale/scale.go:94
};
static_assert(sizeof(ksa_policy_k) == 48, "ksa_policy_k is part of the sweep ABI (48 B)");

// A call replays at most KSA_MAX_CHUNK_TICKS ticks; longer traces are chunked by the caller
// (the carry state makes that exact). Inside a call, ticks are numbered i = 1..T.
#define KSA_MAX_CHUNK_TICKS ((1ll << 30) - 64)
// Cool-downs longer than a chunk behave, inside the chunk, like "blocked until the chunk ends".
#define KSA_COOL_CAP (1 << 30)

// Chunk-local cool-down state of one direction: `s` = the first local tick at which the gate is
// open again, so the gate is a single compare:   blocked at local tick i   <=>   i < s.
// A stamp at local tick i sets s = i + c with c = min(cool_ticks, KSA_COOL_CAP) (so s >= 1 + c);
// the stamp inherited from before the chunk is encoded as
//   s0 = min(max(cool - (tick0 - last), 0), c)            (in [0, c]).
// Equivalence with the reference's `last.Add(cool).After(now)`: SURVEY.md §3.4 gives the tick form
// `k - last < cool` with k = tick0 + i; for an inherited stamp that is i < cool - (tick0 - last),
// and clamping the right-hand side into [0, c] does not change the comparison for
// 1 <= i <= T <= KSA_MAX_CHUNK_TICKS < KSA_COOL_CAP. All values fit int32: a stamp is at most
// i + c <= (2^30 - 64) + 2^30 = 2^31 - 64 (the kernels that number ticks per group of D add at most
// 2D <= 32 to c instead of i, and go down to -(T + 2D) > -2^31).
__device__ __forceinline__ int ksa_local_stamp(long long cool, int c, long long tick0, long long last) {
  long long s = cool - (tick0 - last);
  s = s < 0 ? 0 : s;
  s = s > (long long)c ? (long long)c : s;
  return (int)s;
}
// Stamp to write back: unchanged unless a stamp was taken in this chunk (s > c).
__device__ __forceinline__ long long ksa_global_stamp(int s, int c, long long tick0, long long last) {
  return s > c ? tick0 + (long long)(s - c) : last;
}

struct ksa_trace_regs {
  int r;   // replicas (Deployment.Spec.Replicas)
  int su;  // chunk-local "scale-up gate opens at" tick   (lastScaleUpTime + cool, see above)
  int sd;  // chunk-local "scale-down gate opens at" tick
};

// Per-launch (replay) or per-policy (sweep) constants in the form the step function wants them.
struct ksa_step_consts {
  long long up_msgs, down_msgs;
  int max32, min32, up32, down32;
  int cu, cd;  // min(cool_ticks, KSA_COOL_CAP)
  __device__ __forceinline__ explicit ksa_step_consts(const ksa_policy_k& p)
      : up_msgs(p.up_msgs), down_msgs(p.down_msgs), max32(p.max32), min32(p.min32),
        up32(p.up32), down32(p.down32),
        cu((int)(p.cu > KSA_COOL_CAP ? KSA_COOL_CAP : p.cu)),
        cd((int)(p.cd > KSA_COOL_CAP ? KSA_COOL_CAP : p.cd)) {}
};

// The "wide" event code: event bit k (KSA_EV_*) moved to bit 5k for the six policy events, the two
// failure bits to 30 and 31. The summary kernels build it directly (the same six predicated
// operations that would build the event byte), add it as it is into a packed counter word (six
// 5-bit fields, folded every <= 31 ticks) and feed it to the fingerprint; they never need the
// byte. include/ksa_replay.h documents it as part of the definition of ksa_summary.hash.
#define KSA_W_UP_APPLIED (1u << 0)
#define KSA_W_DOWN_APPLIED (1u << 5)
#define KSA_W_UP_NOOP (1u << 10)
#define KSA_W_DOWN_NOOP (1u << 15)
#define KSA_W_UP_COOL (1u << 20)
#define KSA_W_DOWN_COOL (1u << 25)
#define KSA_W_UP_FAILED (1u << 30)
#define KSA_W_DOWN_FAILED (1u << 31)
__host__ __device__ __forceinline__ unsigned ksa_wide_event(unsigned ev) {
  // three bits at a time: b0 b1 b2 -> bits 0, 5, 10 (times 0x111 lands bit k at k + 4j; the mask
  // keeps j == k; the nine partial products are at distinct positions, so nothing carries)
  const unsigned lo = ((ev & 7u) * 0x111u) & 0x421u;
  const unsigned hi = (((ev >> 3) & 7u) * 0x111u) & 0x421u;
  return lo | (hi << 15) | ((ev >> 6) << 30);
}

// One tick of one trace. `n` = attribute sum, `i` = local tick (1-based). Returns the KSA_EV_* bits.
// Written so that each predicate dies right after it is used: a thread has only 7 predicate
// registers and runs up to 16 traces interleaved.
__device__ __forceinline__ unsigned ksa_step(const ksa_step_consts& c, long long n, int i,
                                             ksa_trace_regs& s) {
  // ---- scale-up half: main.go:51-63 ----
  const bool up_hit = n >= c.up_msgs;                 // main.go:51
  const bool up_go = up_hit && (s.su <= i);           // main.go:52 false -> :57
  const bool up_cool = up_hit && (s.su > i);          // main.go:52 true  -> :53-54 (continue)
  const bool up_applied = up_go && (s.r < c.max32);   // scale.go:62 false -> :66-78
  //    up_noop = up_go && !up_applied                   scale.go:62 true  -> :63-64, returns nil
  unsigned ev = up_cool ? KSA_EV_UP_COOL : 0u;
  ev = up_go ? KSA_EV_UP_NOOP : ev;
  ev = up_applied ? KSA_EV_UP_APPLIED : ev;
  const int r_up = min((int)((unsigned)s.r + (unsigned)c.up32), c.max32);  // scale.go:66-69
  s.r = up_applied ? r_up : s.r;
  s.su = up_go ? i + c.cu : s.su;                     // main.go:62 (also after the no-op)
  // ---- scale-down half: main.go:65-77; skipped when the up half `continue`d ----
  const bool down_hit = (n <= c.down_msgs) && !up_cool;   // main.go:65 (+ :54)
  const bool down_go = down_hit && (s.sd <= i);       // main.go:66 false -> :71
  const bool down_cool = down_hit && (s.sd > i);      // main.go:66 true  -> :67-68
  const bool down_applied = down_go && (s.r > c.min32);   // scale.go:89 (sees the post-up replicas)
  unsigned evd = down_cool ? KSA_EV_DOWN_COOL : 0u;
  evd = down_go ? KSA_EV_DOWN_NOOP : evd;             // scale.go:90-91
  evd = down_applied ? KSA_EV_DOWN_APPLIED : evd;
  const int r_down = max((int)((unsigned)s.r - (unsigned)c.down32), c.min32);  // scale.go:94-97
  s.r = down_applied ? r_down : s.r;
  s.sd = down_go ? i + c.cd : s.sd;                   // main.go:76
  return ev | evd;
}

// The same function with the predication written out in PTX (identical results; the kernels pick
// whichever the variant table says, tools/sweep_variants.py measures all spellings).
__device__ __forceinline__ unsigned ksa_step_ptx(const ksa_step_consts& c, long long n, int i,
                                                 ksa_trace_regs& s) {
  unsigned ev;
  const int icu = i + c.cu, icd = i + c.cd;  // warp-uniform in the replay kernels
  asm("{\n\t"
      ".reg .pred uh, ugo, ucool, uapp, dh, dgo, dcool, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      "setp.ge.s64 uh, %4, %5;\n\t"              // main.go:51
      "setp.le.and.s32 ugo, %2, %11, uh;\n\t"    // main.go:52
      "setp.gt.and.s32 ucool, %2, %11, uh;\n\t"
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"   // scale.go:62
      "@ugo mov.s32 %2, %12;\n\t"                // main.go:62
      "selp.u32 %0, 16, 0, ucool;\n\t"
      "@ugo mov.u32 %0, 4;\n\t"
      "@uapp mov.u32 %0, 1;\n\t"
      "add.s32 t, %1, %9;\n\t"                   // scale.go:66
      "min.s32 t, t, %7;\n\t"                    // scale.go:67-69
      "selp.s32 %1, t, %1, uapp;\n\t"
      "setp.le.and.s64 dh, %4, %6, !ucool;\n\t"  // main.go:65 (+ :54)
      "setp.le.and.s32 dgo, %3, %11, dh;\n\t"    // main.go:66
      "setp.gt.and.s32 dcool, %3, %11, dh;\n\t"
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"   // scale.go:89
      "@dgo mov.s32 %3, %13;\n\t"                // main.go:76
      "@dcool add.u32 %0, %0, 32;\n\t"
      "@dgo add.u32 %0, %0, 8;\n\t"
      "@dapp add.u32 %0, %0, -6;\n\t"
      "sub.s32 u, %1, %10;\n\t"                  // scale.go:94
      "max.s32 u, u, %8;\n\t"                    // scale.go:95-97
      "selp.s32 %1, u, %1, dapp;\n\t"
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(icu), "r"(icd));
  return ev;
}

// Third spelling: as ksa_step_ptx, but every conditional move / add is written as a predicated
// integer multiply-add by `one` (a kernel argument that is always 1 and that ptxas cannot fold).
// On sm_100a IMAD issues on the FMA pipe while ISETP/SEL/VIADDMNMX issue on the half-rate ALU pipe,
// which is the pipe this state machine saturates first (profiles/r01_ncu_full_v14_T2000.txt: ALU
// 56 % active, FMA 12 %). This moves 9 of the 22 ALU operations per decision to the idle pipe, and
// the stamp `s = i * one + cool` needs no precomputed i + cool (the sweep has one cool per policy).
// The spelling the full-trajectory kernels use (event byte out). It differs from ksa_step_fma
// below only in how the cool-down events are entered (from their own "blocked" predicates): with
// 8 traces per thread this order of operations measured 1.01 of the copy bandwidth against 0.99
// for the one below (tools/sweep_variants.py, ids 30 / 32), so the HBM-bound kernels keep it.
__device__ __forceinline__ unsigned ksa_step_fma_byte(const ksa_step_consts& c, long long n, int i, int one,
                                                      ksa_trace_regs& s) {
  unsigned ev;
  asm("{\n\t"
      ".reg .pred uh, ugo, ucool, uapp, dh, dgo, dcool, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      "setp.ge.s64 uh, %4, %5;\n\t"                 // main.go:51
      "setp.le.and.s32 ugo, %2, %11, uh;\n\t"       // main.go:52
      "setp.gt.and.s32 ucool, %2, %11, uh;\n\t"
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"      // scale.go:62
      "@ugo mad.lo.s32 %2, %11, %14, %12;\n\t"      // main.go:62   su = i + cu
      "selp.u32 %0, 16, 0, ucool;\n\t"
      "@ugo mad.lo.s32 %0, %14, 4, %0;\n\t"         // UP_NOOP unless ...
      "@uapp mad.lo.s32 %0, %14, -3, %0;\n\t"       // ... UP_APPLIED
      "add.s32 t, %1, %9;\n\t"                      // scale.go:66
      "min.s32 t, t, %7;\n\t"                       // scale.go:67-69
      "@uapp mad.lo.s32 %1, t, %14, 0;\n\t"
      "setp.le.and.s64 dh, %4, %6, !ucool;\n\t"     // main.go:65 (+ :54)
      "setp.le.and.s32 dgo, %3, %11, dh;\n\t"       // main.go:66
      "setp.gt.and.s32 dcool, %3, %11, dh;\n\t"
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"      // scale.go:89
      "@dgo mad.lo.s32 %3, %11, %14, %13;\n\t"      // main.go:76   sd = i + cd
      "@dcool mad.lo.s32 %0, %14, 32, %0;\n\t"
      "@dgo mad.lo.s32 %0, %14, 8, %0;\n\t"
      "@dapp mad.lo.s32 %0, %14, -6, %0;\n\t"
      "sub.s32 u, %1, %10;\n\t"                     // scale.go:94
      "max.s32 u, u, %8;\n\t"                       // scale.go:95-97
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one));
  return ev;
}

// WIDE = true returns the wide event code (ksa_wide_event of the byte) instead of the byte.
// ORDERED = true is for policies with scale_up_messages > scale_down_messages (the caller checks):
// then the two thresholds cannot both be met in one tick, so the `continue` of main.go:54 never
// has a down branch to skip and the up cool-down predicate is not needed at all (one ISETP less;
// ISETP issues on the half-rate ALU pipe, which is what bounds the summary and sweep kernels).
// The cool-down events are entered from the threshold predicate and corrected when the gate is
// open, so neither direction needs a separate "blocked" compare for its event.
template <bool WIDE = false, bool ORDERED = false>
__device__ __forceinline__ unsigned ksa_step_fma(const ksa_step_consts& c, long long n, int i, int one,
                                                 ksa_trace_regs& s) {
  constexpr int kUpCool = WIDE ? (int)KSA_W_UP_COOL : KSA_EV_UP_COOL;
  constexpr int kUpNoop = WIDE ? (int)KSA_W_UP_NOOP : KSA_EV_UP_NOOP;
  constexpr int kUpApp = WIDE ? (int)KSA_W_UP_APPLIED : KSA_EV_UP_APPLIED;
  constexpr int kDownCool = WIDE ? (int)KSA_W_DOWN_COOL : KSA_EV_DOWN_COOL;
  constexpr int kDownNoop = WIDE ? (int)KSA_W_DOWN_NOOP : KSA_EV_DOWN_NOOP;
  constexpr int kDownApp = WIDE ? (int)KSA_W_DOWN_APPLIED : KSA_EV_DOWN_APPLIED;
  unsigned ev;
#define KSA_STEP_FMA_BODY(DOWN_HIT)                                                              \
  asm("{\n\t"                                                                                    \
      ".reg .pred uh, ugo, ucool, uapp, dh, dgo, dapp;\n\t"                                      \
      ".reg .s32 t, u;\n\t"                                                                      \
      "setp.ge.s64 uh, %4, %5;\n\t"                /* main.go:51 */                              \
      "setp.le.and.s32 ugo, %2, %11, uh;\n\t"      /* main.go:52 false */                        \
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"     /* scale.go:62 */                             \
      "selp.u32 %0, %15, 0, uh;\n\t"               /* UP_COOL unless the gate is open ... */     \
      "@ugo mad.lo.s32 %0, %14, %16, %0;\n\t"      /* ... then UP_NOOP unless ... */             \
      "@uapp mad.lo.s32 %0, %14, %17, %0;\n\t"     /* ... UP_APPLIED */                          \
      "add.s32 t, %1, %9;\n\t"                     /* scale.go:66 */                             \
      "min.s32 t, t, %7;\n\t"                      /* scale.go:67-69 */                          \
      "@uapp mad.lo.s32 %1, t, %14, 0;\n\t"                                                      \
      DOWN_HIT                                     /* main.go:65 (+ :54), reads the old su */    \
      "@ugo mad.lo.s32 %2, %11, %14, %12;\n\t"     /* main.go:62   su = i + cu */                \
      "setp.le.and.s32 dgo, %3, %11, dh;\n\t"      /* main.go:66 false */                        \
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"     /* scale.go:89 */                             \
      "@dgo mad.lo.s32 %3, %11, %14, %13;\n\t"     /* main.go:76   sd = i + cd */                \
      "@dh mad.lo.s32 %0, %14, %18, %0;\n\t"       /* DOWN_COOL unless ... */                    \
      "@dgo mad.lo.s32 %0, %14, %19, %0;\n\t"      /* ... DOWN_NOOP unless ... */                \
      "@dapp mad.lo.s32 %0, %14, %20, %0;\n\t"     /* ... DOWN_APPLIED */                        \
      "sub.s32 u, %1, %10;\n\t"                    /* scale.go:94 */                             \
      "max.s32 u, u, %8;\n\t"                      /* scale.go:95-97 */                          \
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"                                                      \
      "}"                                                                                        \
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)                                              \
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),       \
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one), "n"(kUpCool),                     \
        "n"(kUpNoop - kUpCool), "n"(kUpApp - kUpNoop), "n"(kDownCool), "n"(kDownNoop - kDownCool), \
        "n"(kDownApp - kDownNoop))
  if constexpr (ORDERED)
    KSA_STEP_FMA_BODY("setp.le.s64 dh, %4, %6;\n\t");
  else
    KSA_STEP_FMA_BODY("setp.gt.and.s32 ucool, %2, %11, uh;\n\t"
                      "setp.le.and.s64 dh, %4, %6, !ucool;\n\t");
#undef KSA_STEP_FMA_BODY
  return ev;
}

// The sweep's step when every policy of the launch is ORDERED and has thresholds that fit 32 bits
// (INT_MIN < up <= INT_MAX, INT_MIN <= down < INT_MAX; the host checks). `n32` is the attribute
// sum saturated to int32, computed once per (trace, tick) and shared by all policies of the
// thread; under those bounds  n >= up <=> sat(n) >= up  and  n <= down <=> sat(n) <= down
// (a sum beyond int32 saturates to a value that is still on the same side of both thresholds).
// Two ISETP less than ksa_step_fma<true, true>. Returns the wide event code.
__device__ __forceinline__ unsigned ksa_step_fma_narrow(int up_thr, int down_thr, const ksa_step_consts& c,
                                                        int n32, int i, int one, ksa_trace_regs& s) {
  unsigned ev;
  asm("{\n\t"
      ".reg .pred uh, ugo, uapp, dh, dgo, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      "setp.ge.s32 uh, %4, %5;\n\t"                 // main.go:51
      "setp.le.and.s32 ugo, %2, %11, uh;\n\t"       // main.go:52 false
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"      // scale.go:62
      "@ugo mad.lo.s32 %2, %11, %14, %12;\n\t"      // main.go:62   su = i + cu
      "selp.u32 %0, %15, 0, uh;\n\t"
      "@ugo mad.lo.s32 %0, %14, %16, %0;\n\t"
      "@uapp mad.lo.s32 %0, %14, %17, %0;\n\t"
      "add.s32 t, %1, %9;\n\t"                      // scale.go:66
      "min.s32 t, t, %7;\n\t"                       // scale.go:67-69
      "@uapp mad.lo.s32 %1, t, %14, 0;\n\t"
      "setp.le.s32 dh, %4, %6;\n\t"                 // main.go:65
      "setp.le.and.s32 dgo, %3, %11, dh;\n\t"       // main.go:66 false
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"      // scale.go:89
      "@dgo mad.lo.s32 %3, %11, %14, %13;\n\t"      // main.go:76   sd = i + cd
      "@dh mad.lo.s32 %0, %14, %18, %0;\n\t"
      "@dgo mad.lo.s32 %0, %14, %19, %0;\n\t"
      "@dapp mad.lo.s32 %0, %14, %20, %0;\n\t"
      "sub.s32 u, %1, %10;\n\t"                     // scale.go:94
      "max.s32 u, u, %8;\n\t"                       // scale.go:95-97
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)
      : "r"(n32), "r"(up_thr), "r"(down_thr), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one), "n"((int)KSA_W_UP_COOL),
        "n"((int)KSA_W_UP_NOOP - (int)KSA_W_UP_COOL), "n"((int)KSA_W_UP_APPLIED - (int)KSA_W_UP_NOOP),
        "n"((int)KSA_W_DOWN_COOL), "n"((int)KSA_W_DOWN_NOOP - (int)KSA_W_DOWN_COOL),
        "n"((int)KSA_W_DOWN_APPLIED - (int)KSA_W_DOWN_NOOP));
  return ev;
}
__device__ __forceinline__ int ksa_saturate_s32(long long n) {
  int r;
  asm("cvt.sat.s32.s64 %0, %1;" : "=r"(r) : "l"(n));
  return r;
}

// The sweep's step for a register tile whose policies all have the SAME two thresholds (and are
// narrow, see above): the grid of main.go:86-87 x the other tunables varies the thresholds slowest,
// so the machines of a tile almost always share them. The two threshold tests of main.go:51,65 are
// then made once per (trace, tick) by the caller and reach the machines through values, not
// predicates (a thread has 7 predicate registers for four interleaved machines):
//   iu = n >= up   ? i : INT_MIN        gate test "su <= iu" is false when the threshold is not met
//   id = n <= down ? i : INT_MIN        (stamps are >= 0 in the sweep: ksa_local_stamp(.., 0, 0))
//   e0  = UP_COOL if n >= up, DOWN_COOL if n <= down, else 0     (ordered thresholds: never both)
//   e0n = e0 + (UP_NOOP - UP_COOL)      only selected when the up gate is open
// Per machine: 4 ISETP + 1 SEL + 2 VIADDMNMX on the ALU pipe, 7 IMAD on the FMA pipe (the narrow
// step above: 6 ISETP + 1 SEL + 2 VIADDMNMX + 9 IMAD). Returns the wide event code.
__device__ __forceinline__ unsigned ksa_step_fma_shared(const ksa_step_consts& c, int iu, int id, unsigned e0,
                                                        unsigned e0n, int i, int one, ksa_trace_regs& s) {
  unsigned ev;
  asm("{\n\t"
      ".reg .pred ugo, uapp, dgo, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      "setp.le.s32 ugo, %2, %4;\n\t"                // main.go:51 + :52 false
      "setp.lt.and.s32 uapp, %1, %6, ugo;\n\t"      // scale.go:62
      "selp.u32 %0, %15, %14, ugo;\n\t"             // UP_COOL -> UP_NOOP when the gate is open ...
      "@uapp mad.lo.s32 %0, %13, %16, %0;\n\t"      // ... UP_APPLIED
      "add.s32 t, %1, %8;\n\t"                      // scale.go:66
      "min.s32 t, t, %6;\n\t"                       // scale.go:67-69
      "@uapp mad.lo.s32 %1, t, %13, 0;\n\t"
      "@ugo mad.lo.s32 %2, %10, %13, %11;\n\t"      // main.go:62   su = i + cu
      "setp.le.s32 dgo, %3, %5;\n\t"                // main.go:65 + :66 false
      "setp.gt.and.s32 dapp, %1, %7, dgo;\n\t"      // scale.go:89
      "@dgo mad.lo.s32 %3, %10, %13, %12;\n\t"      // main.go:76   sd = i + cd
      "@dgo mad.lo.s32 %0, %13, %17, %0;\n\t"       // DOWN_COOL -> DOWN_NOOP ...
      "@dapp mad.lo.s32 %0, %13, %18, %0;\n\t"      // ... DOWN_APPLIED
      "sub.s32 u, %1, %9;\n\t"                      // scale.go:94
      "max.s32 u, u, %7;\n\t"                       // scale.go:95-97
      "@dapp mad.lo.s32 %1, u, %13, 0;\n\t"
      "}"
      : "=&r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)
      : "r"(iu), "r"(id), "r"(c.max32), "r"(c.min32), "r"(c.up32), "r"(c.down32), "r"(i), "r"(c.cu),
        "r"(c.cd), "r"(one), "r"(e0), "r"(e0n), "n"((int)KSA_W_UP_APPLIED - (int)KSA_W_UP_NOOP),
        "n"((int)KSA_W_DOWN_NOOP - (int)KSA_W_DOWN_COOL), "n"((int)KSA_W_DOWN_APPLIED - (int)KSA_W_DOWN_NOOP));
  return ev;
}

// The step in a world where calls fail (the *_faults entry points). `f` carries the KSA_FAIL_* bits
// of this (trace, tick). The three error paths of Run all `continue` before any state changes:
//   NumMessages failed  main.go:43-47   nothing else happens in the tick
//   ScaleUp failed      main.go:57-60   no stamp, tick ends (down branch skipped)
//   ScaleDown failed    main.go:71-74   no stamp; an up-scale earlier in the tick stands
// A fail bit whose call is not reached (threshold not met, still cooling down) has no effect.
__device__ __forceinline__ unsigned ksa_step_faults(const ksa_step_consts& c, long long n, unsigned f,
                                                    int i, ksa_trace_regs& s) {
  const bool poll_ok = (f & KSA_FAIL_POLL) == 0;          // main.go:44
  // ---- scale-up half ----
  const bool up_hit = poll_ok && n >= c.up_msgs;          // main.go:51
  const bool up_reach = up_hit && (s.su <= i);            // main.go:52 false: ScaleUp is called
  const bool up_cool = up_hit && (s.su > i);              // main.go:52 true -> :53-54
  const bool up_failed = up_reach && (f & KSA_FAIL_UP);   // main.go:57-60 (continue)
  const bool up_go = up_reach && !(f & KSA_FAIL_UP);      // ScaleUp returned nil
  const bool up_applied = up_go && (s.r < c.max32);       // scale.go:62
  unsigned ev = up_cool ? KSA_EV_UP_COOL : 0u;
  ev = up_go ? KSA_EV_UP_NOOP : ev;
  ev = up_applied ? KSA_EV_UP_APPLIED : ev;
  ev = up_failed ? KSA_EV_UP_FAILED : ev;
  const int r_up = min((int)((unsigned)s.r + (unsigned)c.up32), c.max32);  // scale.go:66-69
  s.r = up_applied ? r_up : s.r;
  s.su = up_go ? i + c.cu : s.su;                         // main.go:62
  // ---- scale-down half: skipped when the up half `continue`d (cool-down or error) ----
  const bool down_hit = poll_ok && (n <= c.down_msgs) && !up_cool && !up_failed;  // main.go:65
  const bool down_reach = down_hit && (s.sd <= i);        // main.go:66 false: ScaleDown is called
  const bool down_cool = down_hit && (s.sd > i);          // main.go:66 true -> :67-68
  const bool down_failed = down_reach && (f & KSA_FAIL_DOWN);  // main.go:71-74
  const bool down_go = down_reach && !(f & KSA_FAIL_DOWN);
  const bool down_applied = down_go && (s.r > c.min32);   // scale.go:89
  unsigned evd = down_cool ? KSA_EV_DOWN_COOL : 0u;
  evd = down_go ? KSA_EV_DOWN_NOOP : evd;
  evd = down_applied ? KSA_EV_DOWN_APPLIED : evd;
  evd = down_failed ? KSA_EV_DOWN_FAILED : evd;
  const int r_down = max((int)((unsigned)s.r - (unsigned)c.down32), c.min32);  // scale.go:94-97
  s.r = down_applied ? r_down : s.r;
  s.sd = down_go ? i + c.cd : s.sd;                       // main.go:76
  return poll_ok ? (ev | evd) : (unsigned)KSA_EV_POLL_FAILED;  // main.go:45
}

// ksa_step_faults with the conditional moves / adds as predicated IMADs (see ksa_step_fma).
// `ucont` = the up half `continue`d (cool-down or failed call) = up_hit && !up_go.
template <bool WIDE = false>
__device__ __forceinline__ unsigned ksa_step_faults_fma(const ksa_step_consts& c, long long n, unsigned f,
                                                        int i, int one, ksa_trace_regs& s) {
  constexpr int kUpCool = WIDE ? (int)KSA_W_UP_COOL : KSA_EV_UP_COOL;
  constexpr int kUpNoop = WIDE ? (int)KSA_W_UP_NOOP : KSA_EV_UP_NOOP;
  constexpr int kUpApp = WIDE ? (int)KSA_W_UP_APPLIED : KSA_EV_UP_APPLIED;
  constexpr int kUpFail = WIDE ? (int)KSA_W_UP_FAILED : KSA_EV_UP_FAILED;
  constexpr int kDownCool = WIDE ? (int)KSA_W_DOWN_COOL : KSA_EV_DOWN_COOL;
  constexpr int kDownNoop = WIDE ? (int)KSA_W_DOWN_NOOP : KSA_EV_DOWN_NOOP;
  constexpr int kDownApp = WIDE ? (int)KSA_W_DOWN_APPLIED : KSA_EV_DOWN_APPLIED;
  constexpr int kDownFail = WIDE ? (int)KSA_W_DOWN_FAILED : KSA_EV_DOWN_FAILED;
  constexpr int kPollFail = WIDE ? (int)(KSA_W_UP_FAILED | KSA_W_DOWN_FAILED) : KSA_EV_POLL_FAILED;
  unsigned ev;
  asm("{\n\t"
      ".reg .pred pok, uh, ureach, ucool, ufail, ugo, uapp, ucont, dh, dreach, dcool, dfail, dgo, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      ".reg .b32 fb;\n\t"
      "and.b32 fb, %15, 1;\n\t"
      "setp.eq.u32 pok, fb, 0;\n\t"                 // main.go:44 err == nil
      "setp.ge.and.s64 uh, %4, %5, pok;\n\t"        // main.go:51
      "setp.le.and.s32 ureach, %2, %11, uh;\n\t"    // main.go:52 false: ScaleUp is called
      "setp.gt.and.s32 ucool, %2, %11, uh;\n\t"     // main.go:52 true
      "and.b32 fb, %15, 2;\n\t"
      "setp.ne.and.u32 ufail, fb, 0, ureach;\n\t"   // main.go:57-60
      "setp.eq.and.u32 ugo, fb, 0, ureach;\n\t"     // ScaleUp returned nil
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"      // scale.go:62
      "@ugo mad.lo.s32 %2, %11, %14, %12;\n\t"      // main.go:62   su = i + cu
      "selp.u32 %0, %16, 0, ucool;\n\t"
      "@ugo mad.lo.s32 %0, %14, %17, %0;\n\t"
      "@uapp mad.lo.s32 %0, %14, %18, %0;\n\t"
      "@ufail mad.lo.s32 %0, %14, %19, %0;\n\t"     // UP_FAILED
      "add.s32 t, %1, %9;\n\t"                      // scale.go:66
      "min.s32 t, t, %7;\n\t"                       // scale.go:67-69
      "@uapp mad.lo.s32 %1, t, %14, 0;\n\t"
      "and.pred ucont, uh, !ugo;\n\t"               // main.go:54 / :59 continue
      "setp.le.and.s64 dh, %4, %6, pok;\n\t"        // main.go:65
      "and.pred dh, dh, !ucont;\n\t"
      "setp.le.and.s32 dreach, %3, %11, dh;\n\t"    // main.go:66 false: ScaleDown is called
      "setp.gt.and.s32 dcool, %3, %11, dh;\n\t"
      "and.b32 fb, %15, 4;\n\t"
      "setp.ne.and.u32 dfail, fb, 0, dreach;\n\t"   // main.go:71-74
      "setp.eq.and.u32 dgo, fb, 0, dreach;\n\t"
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"      // scale.go:89
      "@dgo mad.lo.s32 %3, %11, %14, %13;\n\t"      // main.go:76   sd = i + cd
      "@dcool mad.lo.s32 %0, %14, %20, %0;\n\t"
      "@dgo mad.lo.s32 %0, %14, %21, %0;\n\t"
      "@dapp mad.lo.s32 %0, %14, %22, %0;\n\t"
      "@dfail mad.lo.s32 %0, %14, %23, %0;\n\t"     // DOWN_FAILED
      "sub.s32 u, %1, %10;\n\t"                     // scale.go:94
      "max.s32 u, u, %8;\n\t"                       // scale.go:95-97
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"
      "@!pok mov.u32 %0, %24;\n\t"                  // POLL_FAILED, main.go:45
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one), "r"(f), "n"(kUpCool), "n"(kUpNoop),
        "n"(kUpApp - kUpNoop), "n"(kUpFail), "n"(kDownCool), "n"(kDownNoop), "n"(kDownApp - kDownNoop),
        "n"(kDownFail), "n"(kPollFail));
  return ev;
}

// Second spelling of the step with failures, used by the summary kernels (fewer operations on the
// half-rate ALU pipe, which binds there): the cool-down events are entered from the threshold
// predicates and corrected when the gate is open (as in ksa_step_fma), and a failed call is
// entered arithmetically from its KSA_FAIL_* bit under the "call reached" predicate,
//   W += (f & KSA_FAIL_UP) * (UP_FAILED - UP_COOL) / 2,
// so neither the "blocked" nor the "failed" predicates exist. ORDERED (scale_up_messages >
// scale_down_messages, checked by the caller): both thresholds cannot be met in one tick, so the
// up half never has a down half to skip.
template <bool WIDE, bool ORDERED>
__device__ __forceinline__ unsigned ksa_step_faults_fma2(const ksa_step_consts& c, long long n, unsigned f,
                                                         int i, int one, ksa_trace_regs& s) {
  constexpr int kUpCool = WIDE ? (int)KSA_W_UP_COOL : KSA_EV_UP_COOL;
  constexpr int kUpNoop = WIDE ? (int)KSA_W_UP_NOOP : KSA_EV_UP_NOOP;
  constexpr int kUpApp = WIDE ? (int)KSA_W_UP_APPLIED : KSA_EV_UP_APPLIED;
  constexpr long long kUpFail = WIDE ? (long long)KSA_W_UP_FAILED : KSA_EV_UP_FAILED;
  constexpr int kDownCool = WIDE ? (int)KSA_W_DOWN_COOL : KSA_EV_DOWN_COOL;
  constexpr int kDownNoop = WIDE ? (int)KSA_W_DOWN_NOOP : KSA_EV_DOWN_NOOP;
  constexpr int kDownApp = WIDE ? (int)KSA_W_DOWN_APPLIED : KSA_EV_DOWN_APPLIED;
  constexpr long long kDownFail = WIDE ? (long long)KSA_W_DOWN_FAILED : KSA_EV_DOWN_FAILED;
  constexpr int kPollFail = WIDE ? (int)(KSA_W_UP_FAILED | KSA_W_DOWN_FAILED) : KSA_EV_POLL_FAILED;
  constexpr int kUpFailStep = (int)((kUpFail - kUpCool) / KSA_FAIL_UP);        // times (f & KSA_FAIL_UP)
  constexpr int kDownFailStep = (int)((kDownFail - kDownCool) / KSA_FAIL_DOWN);  // times (f & KSA_FAIL_DOWN)
  static_assert((kUpFail - kUpCool) % KSA_FAIL_UP == 0 && (kDownFail - kDownCool) % KSA_FAIL_DOWN == 0, "");
  unsigned ev;
#define KSA_STEP_FAULTS2_BODY(DOWN_HIT)                                                          \
  asm("{\n\t"                                                                                    \
      ".reg .pred pok, uh, ureach, ugo, uapp, ucont, dh, dreach, dgo, dapp;\n\t"                 \
      ".reg .s32 t, u;\n\t"                                                                      \
      ".reg .b32 fb, fu, fd;\n\t"                                                                \
      "and.b32 fb, %15, 1;\n\t"                                                                  \
      "setp.eq.u32 pok, fb, 0;\n\t"                /* main.go:44 err == nil */                   \
      "setp.ge.and.s64 uh, %4, %5, pok;\n\t"       /* main.go:51 */                              \
      "setp.le.and.s32 ureach, %2, %11, uh;\n\t"   /* main.go:52 false: ScaleUp is called */     \
      "and.b32 fu, %15, 2;\n\t"                                                                  \
      "setp.eq.and.u32 ugo, fu, 0, ureach;\n\t"    /* ScaleUp returned nil */                    \
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"     /* scale.go:62 */                             \
      "selp.u32 %0, %16, 0, uh;\n\t"               /* UP_COOL unless the call is reached ... */  \
      "@ugo mad.lo.s32 %0, %14, %17, %0;\n\t"      /* ... and returns nil: UP_NOOP unless ... */ \
      "@uapp mad.lo.s32 %0, %14, %18, %0;\n\t"     /* ... UP_APPLIED */                          \
      "@ureach mad.lo.s32 %0, fu, %19, %0;\n\t"    /* ... or fails: UP_FAILED, main.go:57-60 */  \
      "add.s32 t, %1, %9;\n\t"                     /* scale.go:66 */                             \
      "min.s32 t, t, %7;\n\t"                      /* scale.go:67-69 */                          \
      "@uapp mad.lo.s32 %1, t, %14, 0;\n\t"                                                      \
      DOWN_HIT                                     /* main.go:65, after :54 / :59 continue */    \
      "@ugo mad.lo.s32 %2, %11, %14, %12;\n\t"     /* main.go:62   su = i + cu */                \
      "setp.le.and.s32 dreach, %3, %11, dh;\n\t"   /* main.go:66 false: ScaleDown is called */   \
      "and.b32 fd, %15, 4;\n\t"                                                                  \
      "setp.eq.and.u32 dgo, fd, 0, dreach;\n\t"                                                  \
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"     /* scale.go:89 */                             \
      "@dgo mad.lo.s32 %3, %11, %14, %13;\n\t"     /* main.go:76   sd = i + cd */                \
      "@dh mad.lo.s32 %0, %14, %20, %0;\n\t"       /* DOWN_COOL unless ... */                    \
      "@dgo mad.lo.s32 %0, %14, %21, %0;\n\t"      /* ... DOWN_NOOP unless ... */                \
      "@dapp mad.lo.s32 %0, %14, %22, %0;\n\t"     /* ... DOWN_APPLIED */                        \
      "@dreach mad.lo.s32 %0, fd, %23, %0;\n\t"    /* ... or DOWN_FAILED, main.go:71-74 */       \
      "sub.s32 u, %1, %10;\n\t"                    /* scale.go:94 */                             \
      "max.s32 u, u, %8;\n\t"                      /* scale.go:95-97 */                          \
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"                                                      \
      "@!pok mov.u32 %0, %24;\n\t"                 /* POLL_FAILED, main.go:45 */                 \
      "}"                                                                                        \
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)                                              \
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),       \
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one), "r"(f), "n"(kUpCool),             \
        "n"(kUpNoop - kUpCool), "n"(kUpApp - kUpNoop), "n"(kUpFailStep), "n"(kDownCool),         \
        "n"(kDownNoop - kDownCool), "n"(kDownApp - kDownNoop), "n"(kDownFailStep), "n"(kPollFail))
  if constexpr (ORDERED)
    KSA_STEP_FAULTS2_BODY("setp.le.and.s64 dh, %4, %6, pok;\n\t");
  else
    KSA_STEP_FAULTS2_BODY("and.pred ucont, uh, !ugo;\n\t"
                          "setp.le.and.s64 dh, %4, %6, pok;\n\t"
                          "and.pred dh, dh, !ucont;\n\t");
#undef KSA_STEP_FAULTS2_BODY
  return ev;
}

// Third spelling, for the summary kernels (wide event code only). Against ksa_step_faults_fma2:
//  - the three failure counters of the summary are accumulated here, straight from the mask bits
//    under the "call reached" predicates, into one packed word `pkf` (KSA_PKF_* below): three
//    IMADs instead of the six ALU operations that recovered them from the event code afterwards;
//  - a failed poll enters the event code as (f & KSA_FAIL_POLL) * POLL_FAILED up front (both
//    thresholds are then unmet, so nothing else is added) instead of a final select.
// pkf fields (10 bits each, flushed with the other second-level counters, ksa_summary_flush_faults3):
//   up_failed   in units of KSA_FAIL_UP (2)   at bit 0   -> bits 1..10
//   down_failed in units of KSA_FAIL_DOWN (4) at bit 10  -> bits 12..21
//   poll_failed in units of 1                 at bit 22  -> bits 22..31
#define KSA_PKF_UP_SHIFT 1
#define KSA_PKF_DOWN_SHIFT 12
#define KSA_PKF_POLL_SHIFT 22
template <bool ORDERED>
__device__ __forceinline__ unsigned ksa_step_faults_fma3(const ksa_step_consts& c, long long n, unsigned f,
                                                         int i, int one, ksa_trace_regs& s, unsigned& pkf) {
  constexpr int kUpFailStep = (int)(((long long)KSA_W_UP_FAILED - (long long)KSA_W_UP_COOL) / KSA_FAIL_UP);
  constexpr int kDownFailStep = (int)(((long long)KSA_W_DOWN_FAILED - (long long)KSA_W_DOWN_COOL) / KSA_FAIL_DOWN);
  unsigned ev;
#define KSA_STEP_FAULTS3_BODY(DOWN_HIT)                                                          \
  asm("{\n\t"                                                                                    \
      ".reg .pred pok, uh, ureach, ugo, uapp, ucont, dh, dreach, dgo, dapp;\n\t"                 \
      ".reg .s32 t, u;\n\t"                                                                      \
      ".reg .b32 fb, fu, fd;\n\t"                                                                \
      "and.b32 fb, %16, 1;\n\t"                                                                  \
      "setp.eq.u32 pok, fb, 0;\n\t"                /* main.go:44 err == nil */                   \
      "mad.lo.s32 %0, fb, %24, 0;\n\t"             /* POLL_FAILED, main.go:45 */                 \
      "mad.lo.u32 %4, fb, %25, %4;\n\t"            /* count it */                                \
      "setp.ge.and.s64 uh, %5, %6, pok;\n\t"       /* main.go:51 */                              \
      "setp.le.and.s32 ureach, %2, %12, uh;\n\t"   /* main.go:52 false: ScaleUp is called */     \
      "and.b32 fu, %16, 2;\n\t"                                                                  \
      "setp.eq.and.u32 ugo, fu, 0, ureach;\n\t"    /* ScaleUp returned nil */                    \
      "setp.lt.and.s32 uapp, %1, %8, ugo;\n\t"     /* scale.go:62 */                             \
      "@uh mad.lo.s32 %0, %15, %17, %0;\n\t"       /* UP_COOL unless the call is reached ... */  \
      "@ugo mad.lo.s32 %0, %15, %18, %0;\n\t"      /* ... and returns nil: UP_NOOP unless ... */ \
      "@uapp mad.lo.s32 %0, %15, %19, %0;\n\t"     /* ... UP_APPLIED */                          \
      "@ureach mad.lo.s32 %0, fu, %20, %0;\n\t"    /* ... or fails: UP_FAILED, main.go:57-60 */  \
      "@ureach mad.lo.u32 %4, fu, 1, %4;\n\t"      /* count it (units of 2) */                   \
      "add.s32 t, %1, %10;\n\t"                    /* scale.go:66 */                             \
      "min.s32 t, t, %8;\n\t"                      /* scale.go:67-69 */                          \
      "@uapp mad.lo.s32 %1, t, %15, 0;\n\t"                                                      \
      DOWN_HIT                                     /* main.go:65, after :54 / :59 continue */    \
      "@ugo mad.lo.s32 %2, %12, %15, %13;\n\t"     /* main.go:62   su = i + cu */                \
      "setp.le.and.s32 dreach, %3, %12, dh;\n\t"   /* main.go:66 false: ScaleDown is called */   \
      "and.b32 fd, %16, 4;\n\t"                                                                  \
      "setp.eq.and.u32 dgo, fd, 0, dreach;\n\t"                                                  \
      "setp.gt.and.s32 dapp, %1, %9, dgo;\n\t"     /* scale.go:89 */                             \
      "@dgo mad.lo.s32 %3, %12, %15, %14;\n\t"     /* main.go:76   sd = i + cd */                \
      "@dh mad.lo.s32 %0, %15, %21, %0;\n\t"       /* DOWN_COOL unless ... */                    \
      "@dgo mad.lo.s32 %0, %15, %22, %0;\n\t"      /* ... DOWN_NOOP unless ... */                \
      "@dapp mad.lo.s32 %0, %15, %23, %0;\n\t"     /* ... DOWN_APPLIED */                        \
      "@dreach mad.lo.s32 %0, fd, %26, %0;\n\t"    /* ... or DOWN_FAILED, main.go:71-74 */       \
      "@dreach mad.lo.u32 %4, fd, 1024, %4;\n\t"   /* count it (units of 4, at bit 10) */        \
      "sub.s32 u, %1, %11;\n\t"                    /* scale.go:94 */                             \
      "max.s32 u, u, %9;\n\t"                      /* scale.go:95-97 */                          \
      "@dapp mad.lo.s32 %1, u, %15, 0;\n\t"                                                      \
      "}"                                                                                        \
      : "=&r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd), "+r"(pkf)                                  \
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),       \
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one), "r"(f), "n"((int)KSA_W_UP_COOL),  \
        "n"((int)KSA_W_UP_NOOP - (int)KSA_W_UP_COOL), "n"((int)KSA_W_UP_APPLIED - (int)KSA_W_UP_NOOP), \
        "n"(kUpFailStep), "n"((int)KSA_W_DOWN_COOL), "n"((int)KSA_W_DOWN_NOOP - (int)KSA_W_DOWN_COOL), \
        "n"((int)KSA_W_DOWN_APPLIED - (int)KSA_W_DOWN_NOOP),                                     \
        "n"((int)(KSA_W_UP_FAILED | KSA_W_DOWN_FAILED)), "n"(1 << KSA_PKF_POLL_SHIFT), "n"(kDownFailStep))
  if constexpr (ORDERED)
    KSA_STEP_FAULTS3_BODY("setp.le.and.s64 dh, %5, %7, pok;\n\t");
  else
    KSA_STEP_FAULTS3_BODY("and.pred ucont, uh, !ugo;\n\t"
                          "setp.le.and.s64 dh, %5, %7, pok;\n\t"
                          "and.pred dh, dh, !ucont;\n\t");
#undef KSA_STEP_FAULTS3_BODY
  return ev;
}

// The rare path behind ksa_summary_acc::bad: replay one trace again, plainly, and return the exact
// 64-bit sum of its T post-tick replica counts. `depth` / `fail` point at the trace's own column
// (element (t, a) at depth[t * row + a * ld], fail[t * fld]; fail may be NULL); `s` is the state the
// trace started the call with.
static __device__ __noinline__ long long ksa_exact_sum(const ksa_policy_k pk, const long long* depth, long long row,
                                                       long long ld, int A, const unsigned char* fail,
                                                       long long fld, int T, ksa_trace_regs s) {
  const ksa_step_consts c(pk);
  unsigned long long acc = 0ull;
  for (int t = 0; t < T; ++t) {
    unsigned long long n = 0ull;
    for (int a = 0; a < A; ++a) n += (unsigned long long)depth[(long long)t * row + a * ld];
    const unsigned f = fail ? (unsigned)fail[(long long)t * fld] : 0u;
    ksa_step_faults(c, (long long)n, f, t + 1, s);  // with f == 0 this is ksa_step
    acc += (unsigned long long)(long long)s.r;
  }
  return (long long)acc;
}

// d = a * b + c as a true IMAD (FMA pipe): ptxas cannot fold it when b is not a compile-time
// constant, which is what the `one` kernel argument is for.
__device__ __forceinline__ unsigned ksa_mad_u32(unsigned a, unsigned b, unsigned c) {
  unsigned d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ long long ksa_mad_wide_s32(int a, int b, long long c) {
  long long d;
  asm("mad.wide.s32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
  return d;
}

// Per-trace accumulators of the summary kernels (registers). The six event counters are kept as
// 5-bit fields of one register, in the layout of the wide event code, so that a tick's code is
// simply added; they are folded into 10-bit fields before a field can overflow (every <= 31 ticks,
// ksa_summary_fold) and those into the 32-bit counters of the record (ksa_summary_flush).
// The sum of the replica counts is kept the same way: `sw` takes one 32-bit multiply-add per tick
// and is folded into the 64-bit `sum` with the counters. A window of <= 31 ticks cannot overflow
// 32 bits while every replica count is inside +-KSA_SUM_SAFE (31 * 2^26 < 2^31); each fold checks
// the running minimum / maximum (which cover the window) and sets `bad` otherwise, and a trace
// with `bad` set gets its sum recomputed exactly at the end of the kernel (ksa_exact_sum; only
// deployments with more than 67 million replicas ever go there). A 64-bit add per tick costs
// three instructions (IMAD.WIDE + IADD3 + IMAD.X in SASS) against one.
struct ksa_summary_acc {
  int rmin, rmax;
  int rstash;         // first replica count of a pair of ticks (KSA_TICK_FIRST), see ksa_summary_tick
  unsigned pk;        // six 5-bit event counters, field k at bit 5k (bits 30-31: don't care)
  unsigned lva, lvb;  // second level, 10-bit fields: lva = events 0, 2, 4 at bits 0, 10, 20; lvb = 1, 3, 5
  unsigned pkf;       // *_faults kernels only, 10-bit fields (KSA_PKF_* or the older layout, per step)
  int sw;             // sum of the replica counts since the last fold
  unsigned bad;       // != 0: a window may have overflowed `sw`; the sum is recomputed at the end
  long long sum;
  unsigned ha, hb;    // the two lanes of the fingerprint
};
#define KSA_SUM_SAFE (1 << 26)
// pk is folded into lva/lvb after at most KSA_FOLD_TICKS ticks, those are flushed to the 32-bit
// counters of the record after at most KSA_FLUSH_FOLDS folds (30 * 31 = 930, plus the at most 93 ticks
// that the end of a kernel adds before its last flush: <= 1023).
#define KSA_FOLD_TICKS 31
#define KSA_FLUSH_FOLDS 30

// Trajectory fingerprint (ksa_summary.hash, include/ksa_replay.h): two 32-bit lanes,
//   lo <- lo * KSA_HASH_MUL_LO + (uint32)replicas      hi <- hi * KSA_HASH_MUL_HI + wide_event
// one IMAD each; hash = hi << 32 | lo.
#define KSA_HASH_MUL_LO 0x9E3779B1u
#define KSA_HASH_MUL_HI 0x85EBCA6Bu
__host__ __device__ __forceinline__ unsigned long long ksa_hash_step(unsigned long long h, int r, unsigned ev) {
  const unsigned lo = (unsigned)h * KSA_HASH_MUL_LO + (unsigned)r;
  const unsigned hi = (unsigned)(h >> 32) * KSA_HASH_MUL_HI + ksa_wide_event(ev);
  return ((unsigned long long)hi << 32) | lo;
}

// One tick of summary accumulation; `w` = the wide event code of the tick. FMA = true spells the
// adds as IMADs by `one` (FMA pipe). `phase` (a compile-time constant after unrolling) lets two
// consecutive ticks share their min / max update: sm_100a has three-input VIMNMX3, so
// min(min(rmin, r_first), r_second) is one instruction instead of two. KSA_TICK_FIRST only stashes
// its replica count; the caller guarantees that a KSA_TICK_SECOND follows before anything reads
// rmin / rmax (or calls ksa_summary_minmax_finish when the last tick of a trace is a FIRST).
enum { KSA_TICK_SINGLE = 0, KSA_TICK_FIRST = 1, KSA_TICK_SECOND = 2 };
template <bool FMA>
__device__ __forceinline__ void ksa_summary_tick(ksa_summary_acc& m, int r, unsigned w, int one,
                                                 int phase = KSA_TICK_SINGLE) {
  if (phase == KSA_TICK_FIRST) {
    m.rstash = r;
  } else if (phase == KSA_TICK_SECOND) {
    m.rmin = min(min(m.rmin, m.rstash), r);
    m.rmax = max(max(m.rmax, m.rstash), r);
  } else {
    m.rmin = min(m.rmin, r);
    m.rmax = max(m.rmax, r);
  }
  if constexpr (FMA) {
    m.pk = ksa_mad_u32(w, (unsigned)one, m.pk);
    m.sw = (int)ksa_mad_u32((unsigned)r, (unsigned)one, (unsigned)m.sw);
  } else {
    m.pk += w;
    m.sw = (int)((unsigned)m.sw + (unsigned)r);
  }
  m.ha = m.ha * KSA_HASH_MUL_LO + (unsigned)r;
  m.hb = m.hb * KSA_HASH_MUL_HI + w;
}
__device__ __forceinline__ void ksa_summary_minmax_finish(ksa_summary_acc& m) {
  m.rmin = min(m.rmin, m.rstash);
  m.rmax = max(m.rmax, m.rstash);
}

// Failure counters of the *_faults kernels. `f` is the KSA_FAIL_* input byte: a poll fails exactly
// when its bit is set, and a failed poll shows as both failure bits in the event code.
__device__ __forceinline__ void ksa_summary_tick_faults(ksa_summary_acc& m, unsigned w, unsigned f) {
  m.pkf += (((w >> 30) * 0x201u) & 0x401u) + ((f & KSA_FAIL_POLL) << 20);
}

// First level: move the 5-bit fields into the 10-bit ones and the window sum into the 64-bit one.
// Callers fold only when no KSA_TICK_FIRST is pending, so rmin / rmax cover every tick of the window.
__device__ __forceinline__ void ksa_summary_fold(ksa_summary_acc& m) {
  m.lva += m.pk & 0x01F07C1Fu;
  m.lvb += (m.pk >> 5) & 0x01F07C1Fu;
  m.pk = 0u;
  m.sum = (long long)((unsigned long long)m.sum + (unsigned long long)(long long)m.sw);
  m.sw = 0;
  m.bad |= (unsigned)(m.rmin < -KSA_SUM_SAFE) | (unsigned)(m.rmax >= KSA_SUM_SAFE);
}
__device__ __forceinline__ void ksa_summary_acc_init(ksa_summary_acc& m, int rmin, int rmax, long long sum,
                                                     unsigned long long hash) {
  m.rmin = rmin;
  m.rmax = rmax;
  m.rstash = 0;
  m.pk = m.lva = m.lvb = m.pkf = 0u;
  m.sw = 0;
  m.bad = 0u;
  m.sum = sum;
  m.ha = (unsigned)hash;
  m.hb = (unsigned)(hash >> 32);
}

// Second level: fold, then add the 10-bit fields into the 32-bit counters of a ksa_summary record.
__device__ __forceinline__ void ksa_summary_flush(ksa_summary_acc& m, ksa_summary* q) {
  ksa_summary_fold(m);
  q->up_applied += m.lva & 0x3ffu;
  q->up_noop += (m.lva >> 10) & 0x3ffu;
  q->up_cool += (m.lva >> 20) & 0x3ffu;
  q->down_applied += m.lvb & 0x3ffu;
  q->down_noop += (m.lvb >> 10) & 0x3ffu;
  q->down_cool += (m.lvb >> 20) & 0x3ffu;
  m.lva = m.lvb = 0u;
}
// The same for the packed word ksa_step_faults_fma3 accumulates (KSA_PKF_*): the three counts are
// already separate (a failed poll reaches neither call).
__device__ __forceinline__ void ksa_summary_flush_faults3(ksa_summary_acc& m, ksa_summary* q) {
  q->up_failed += (m.pkf >> KSA_PKF_UP_SHIFT) & 0x3ffu;
  q->down_failed += (m.pkf >> KSA_PKF_DOWN_SHIFT) & 0x3ffu;
  q->poll_failed += (m.pkf >> KSA_PKF_POLL_SHIFT) & 0x3ffu;
  m.pkf = 0u;
}
__device__ __forceinline__ void ksa_summary_flush_faults(ksa_summary_acc& m, ksa_summary* q) {
  const unsigned polls = (m.pkf >> 20) & 0x3ffu;
  q->poll_failed += polls;
  q->up_failed += (m.pkf & 0x3ffu) - polls;
  q->down_failed += ((m.pkf >> 10) & 0x3ffu) - polls;
  m.pkf = 0u;
}
