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
#define KSA_MAX_CHUNK_TICKS (1ll << 30)
// Cool-downs longer than a chunk behave, inside the chunk, like "blocked until the chunk ends".
#define KSA_COOL_CAP ((1 << 30) + 1)

// Chunk-local cool-down state of one direction: `s` = the first local tick at which the gate is
// open again, so the gate is a single compare:   blocked at local tick i   <=>   i < s.
// A stamp at local tick i sets s = i + c with c = min(cool_ticks, KSA_COOL_CAP) (so s >= 1 + c);
// the stamp inherited from before the chunk is encoded as
//   s0 = min(max(cool - (tick0 - last), 0), c)            (in [0, c]).
// Equivalence with the reference's `last.Add(cool).After(now)`: SURVEY.md §3.4 gives the tick form
// `k - last < cool` with k = tick0 + i; for an inherited stamp that is i < cool - (tick0 - last),
// and clamping the right-hand side into [0, c] does not change the comparison for
// 1 <= i <= T <= 2^30 < KSA_COOL_CAP. All values stay below 2^31: i + c <= 2^30 + 2^30 + 1.
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
__device__ __forceinline__ unsigned ksa_step_fma(const ksa_step_consts& c, long long n, int i, int one,
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
__device__ __forceinline__ unsigned ksa_step_faults_fma(const ksa_step_consts& c, long long n, unsigned f,
                                                        int i, int one, ksa_trace_regs& s) {
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
      "selp.u32 %0, 16, 0, ucool;\n\t"
      "@ugo mad.lo.s32 %0, %14, 4, %0;\n\t"
      "@uapp mad.lo.s32 %0, %14, -3, %0;\n\t"
      "@ufail mad.lo.s32 %0, %14, 64, %0;\n\t"      // KSA_EV_UP_FAILED
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
      "@dcool mad.lo.s32 %0, %14, 32, %0;\n\t"
      "@dgo mad.lo.s32 %0, %14, 8, %0;\n\t"
      "@dapp mad.lo.s32 %0, %14, -6, %0;\n\t"
      "@dfail mad.lo.s32 %0, %14, 128, %0;\n\t"     // KSA_EV_DOWN_FAILED
      "sub.s32 u, %1, %10;\n\t"                     // scale.go:94
      "max.s32 u, u, %8;\n\t"                       // scale.go:95-97
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"
      "@!pok mov.u32 %0, 192;\n\t"                  // KSA_EV_POLL_FAILED, main.go:45
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.su), "+r"(s.sd)
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(c.cu), "r"(c.cd), "r"(one), "r"(f));
  return ev;
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
// 8-bit fields of two registers (pk4: bits 0..3 of the event byte, one field per byte; pk2: bits 4
// and 5 at bit 8 and bit 16) and must be added into 32-bit counters before a field can overflow,
// i.e. at least every KSA_FLUSH_TICKS ticks (ksa_summary_flush).
struct ksa_summary_acc {
  int rmin, rmax;
  unsigned pk4, pk2;
  unsigned pkf;  // *_faults kernels only: UP_FAILED bit at field 0, DOWN_FAILED at 8, failed polls at 16
  long long sum;
  unsigned long long hash;
};
#define KSA_FLUSH_TICKS 255

// Trajectory fingerprint step (ksa_summary.hash, include/ksa_replay.h):
//   h <- h * M + ((uint64)event << 32 | (uint32)replicas)
// Three IMADs on the FMA pipe; the addend needs no shifting or masking.
#define KSA_HASH_MUL 0x9E3779B97F4A7C15ull
__device__ __forceinline__ unsigned long long ksa_hash_step(unsigned long long h, int r, unsigned ev) {
  return h * KSA_HASH_MUL + (((unsigned long long)ev << 32) | (unsigned long long)(unsigned)r);
}

// One tick of summary accumulation. FMA = true spells the adds as IMADs by `one` (FMA pipe).
// `ev6` = the low six event bits (the *_faults kernels mask the two failure bits off; everywhere
// else ev6 == ev).
template <bool FMA>
__device__ __forceinline__ void ksa_summary_tick(ksa_summary_acc& m, int r, unsigned ev, unsigned ev6,
                                                 int one) {
  m.rmin = min(m.rmin, r);
  m.rmax = max(m.rmax, r);
  // spread event bits 0..3 into four byte fields, bits 4..5 into two: one IMAD + one LOP3 each.
  // ev6 < 64, so the shifted copies (shifts 0/7/14/21 and 4/11) never overlap: no carries.
  const unsigned f4 = (ev6 * 0x00204081u) & 0x01010101u;
  const unsigned f2 = (ev6 * 0x00000810u) & 0x00010100u;
  if constexpr (FMA) {
    m.pk4 = ksa_mad_u32(f4, (unsigned)one, m.pk4);
    m.pk2 = ksa_mad_u32(f2, (unsigned)one, m.pk2);
    m.sum = ksa_mad_wide_s32(r, one, m.sum);
  } else {
    m.pk4 += f4;
    m.pk2 += f2;
    m.sum = (long long)((unsigned long long)m.sum + (unsigned long long)(long long)r);
  }
  m.hash = ksa_hash_step(m.hash, r, ev);
}

// Failure counters of the *_faults kernels. `f` is the KSA_FAIL_* input byte: a poll fails exactly
// when its bit is set, and a failed poll shows as both failure bits in the event byte.
__device__ __forceinline__ void ksa_summary_tick_faults(ksa_summary_acc& m, unsigned ev, unsigned f) {
  m.pkf += (((ev >> 6) * 0x81u) & 0x101u) + ((f & KSA_FAIL_POLL) << 16);
}
__device__ __forceinline__ void ksa_summary_flush_faults(ksa_summary_acc& m, ksa_summary* q) {
  const unsigned polls = (m.pkf >> 16) & 0xffu;
  q->poll_failed += polls;
  q->up_failed += (m.pkf & 0xffu) - polls;
  q->down_failed += ((m.pkf >> 8) & 0xffu) - polls;
  m.pkf = 0u;
}

// Add the packed fields into the 32-bit counters of a ksa_summary record and clear them.
__device__ __forceinline__ void ksa_summary_flush(ksa_summary_acc& m, ksa_summary* q) {
  q->up_applied += m.pk4 & 0xffu;
  q->down_applied += (m.pk4 >> 8) & 0xffu;
  q->up_noop += (m.pk4 >> 16) & 0xffu;
  q->down_noop += m.pk4 >> 24;
  q->up_cool += (m.pk2 >> 8) & 0xffu;
  q->down_cool += (m.pk2 >> 16) & 0xffu;
  m.pk4 = m.pk2 = 0u;
}
