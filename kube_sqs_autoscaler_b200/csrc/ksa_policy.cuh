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

// Chunk-local cool-down state of one direction: `l` = local tick of the last stamp, shifted so
// that the gate is a single compare:   blocked at local tick i   <=>   i - c < l,
// with c = min(cool_ticks, KSA_COOL_CAP). A stamp at local tick i sets l = i (>= 1); values <= 0
// encode the stamp inherited from before the chunk:
//   l0 = min(max(cool - (tick0 - last), 0), c) - c         (in [-c, 0]).
// Proof of equivalence with `k - last < cool` (k = tick0 + i): SURVEY.md §3.4 gives the tick
// form; for an inherited stamp, blocked <=> i < cool - (tick0 - last) =: s; clamping s into
// [0, c] does not change `i < s` for 1 <= i <= T <= 2^30 < c when cool > c, and is the identity
// otherwise.
__device__ __forceinline__ int ksa_local_stamp(long long cool, int c, long long tick0, long long last) {
  long long s = cool - (tick0 - last);
  s = s < 0 ? 0 : s;
  s = s > (long long)c ? (long long)c : s;
  return (int)s - c;
}
// Stamp to write back: unchanged unless a stamp was taken in this chunk (l >= 1).
__device__ __forceinline__ long long ksa_global_stamp(int l, long long tick0, long long last) {
  return l >= 1 ? tick0 + (long long)l : last;
}

struct ksa_trace_regs {
  int r;   // replicas (Deployment.Spec.Replicas)
  int lu;  // chunk-local lastScaleUpTime   (see above)
  int ld;  // chunk-local lastScaleDownTime
};

// Per-launch constants in the form the step function wants them.
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

// One tick of one trace. `n` = attribute sum, `i` = local tick (1-based), `iu` = i - cu,
// `id` = i - cd (warp-uniform, hoisted by the caller). Returns the KSA_EV_* bits.
// Written so that each predicate dies right after it is used: a thread has only 7 predicate
// registers and runs 2-8 traces interleaved.
__device__ __forceinline__ unsigned ksa_step(const ksa_step_consts& c, long long n, int i, int iu,
                                             int id, ksa_trace_regs& s) {
  // ---- scale-up half: main.go:51-63 ----
  const bool up_hit = n >= c.up_msgs;                 // main.go:51
  const bool up_go = up_hit && (s.lu <= iu);          // main.go:52 false -> :57
  const bool up_cool = up_hit && (s.lu > iu);         // main.go:52 true  -> :53-54 (continue)
  const bool up_applied = up_go && (s.r < c.max32);   // scale.go:62 false -> :66-78
  //    up_noop = up_go && !up_applied                   scale.go:62 true  -> :63-64, returns nil
  unsigned ev = up_cool ? KSA_EV_UP_COOL : 0u;
  ev = up_go ? KSA_EV_UP_NOOP : ev;
  ev = up_applied ? KSA_EV_UP_APPLIED : ev;
  const int r_up = min((int)((unsigned)s.r + (unsigned)c.up32), c.max32);  // scale.go:66-69
  s.r = up_applied ? r_up : s.r;
  s.lu = up_go ? i : s.lu;                            // main.go:62 (also after the no-op)
  // ---- scale-down half: main.go:65-77; skipped when the up half `continue`d ----
  const bool down_hit = (n <= c.down_msgs) && !up_cool;   // main.go:65 (+ :54)
  const bool down_go = down_hit && (s.ld <= id);      // main.go:66 false -> :71
  const bool down_cool = down_hit && (s.ld > id);     // main.go:66 true  -> :67-68
  const bool down_applied = down_go && (s.r > c.min32);   // scale.go:89 (sees the post-up replicas)
  unsigned evd = down_cool ? KSA_EV_DOWN_COOL : 0u;
  evd = down_go ? KSA_EV_DOWN_NOOP : evd;             // scale.go:90-91
  evd = down_applied ? KSA_EV_DOWN_APPLIED : evd;
  const int r_down = max((int)((unsigned)s.r - (unsigned)c.down32), c.min32);  // scale.go:94-97
  s.r = down_applied ? r_down : s.r;
  s.ld = down_go ? i : s.ld;                          // main.go:76
  return ev | evd;
}

// The same function with the predication written out in PTX (identical results; the kernels pick
// whichever the variant table says, tools/sweep_variants.py measures both).
__device__ __forceinline__ unsigned ksa_step_ptx(const ksa_step_consts& c, long long n, int i, int iu,
                                                 int id, ksa_trace_regs& s) {
  unsigned ev;
  asm("{\n\t"
      ".reg .pred uh, ugo, ucool, uapp, dh, dgo, dcool, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      "setp.ge.s64 uh, %4, %5;\n\t"              // main.go:51
      "setp.le.and.s32 ugo, %2, %12, uh;\n\t"    // main.go:52
      "setp.gt.and.s32 ucool, %2, %12, uh;\n\t"
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"   // scale.go:62
      "@ugo mov.s32 %2, %11;\n\t"                // main.go:62
      "selp.u32 %0, 16, 0, ucool;\n\t"
      "@ugo mov.u32 %0, 4;\n\t"
      "@uapp mov.u32 %0, 1;\n\t"
      "add.s32 t, %1, %9;\n\t"                   // scale.go:66
      "min.s32 t, t, %7;\n\t"                    // scale.go:67-69
      "selp.s32 %1, t, %1, uapp;\n\t"
      "setp.le.and.s64 dh, %4, %6, !ucool;\n\t"  // main.go:65 (+ :54)
      "setp.le.and.s32 dgo, %3, %13, dh;\n\t"    // main.go:66
      "setp.gt.and.s32 dcool, %3, %13, dh;\n\t"
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"   // scale.go:89
      "@dgo mov.s32 %3, %11;\n\t"                // main.go:76
      "@dcool add.u32 %0, %0, 32;\n\t"
      "@dgo add.u32 %0, %0, 8;\n\t"
      "@dapp add.u32 %0, %0, -6;\n\t"
      "sub.s32 u, %1, %10;\n\t"                  // scale.go:94
      "max.s32 u, u, %8;\n\t"                    // scale.go:95-97
      "selp.s32 %1, u, %1, dapp;\n\t"
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.lu), "+r"(s.ld)
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(iu), "r"(id));
  return ev;
}

// Third spelling: as ksa_step_ptx, but every conditional move / add is written as a predicated
// integer multiply-add by `one` (a kernel argument that is always 1 and that ptxas cannot fold).
// On sm_100a IMAD issues on the FMA pipe while ISETP/SEL/VIADDMNMX issue on the half-rate ALU pipe,
// which is the pipe this state machine saturates first (profiles/r01_ncu_full_v14_T2000.txt: ALU
// 56 % active, FMA 12 %). This moves 9 of the 22 ALU operations per decision to the idle pipe.
__device__ __forceinline__ unsigned ksa_step_fma(const ksa_step_consts& c, long long n, int i, int iu,
                                                 int id, int one, ksa_trace_regs& s) {
  unsigned ev;
  asm("{\n\t"
      ".reg .pred uh, ugo, ucool, uapp, dh, dgo, dcool, dapp;\n\t"
      ".reg .s32 t, u;\n\t"
      "setp.ge.s64 uh, %4, %5;\n\t"                 // main.go:51
      "setp.le.and.s32 ugo, %2, %12, uh;\n\t"       // main.go:52
      "setp.gt.and.s32 ucool, %2, %12, uh;\n\t"
      "setp.lt.and.s32 uapp, %1, %7, ugo;\n\t"      // scale.go:62
      "@ugo mad.lo.s32 %2, %11, %14, 0;\n\t"        // main.go:62   lu = i
      "selp.u32 %0, 16, 0, ucool;\n\t"
      "@ugo mad.lo.s32 %0, %14, 4, %0;\n\t"         // UP_NOOP unless ...
      "@uapp mad.lo.s32 %0, %14, -3, %0;\n\t"       // ... UP_APPLIED
      "add.s32 t, %1, %9;\n\t"                      // scale.go:66
      "min.s32 t, t, %7;\n\t"                       // scale.go:67-69
      "@uapp mad.lo.s32 %1, t, %14, 0;\n\t"
      "setp.le.and.s64 dh, %4, %6, !ucool;\n\t"     // main.go:65 (+ :54)
      "setp.le.and.s32 dgo, %3, %13, dh;\n\t"       // main.go:66
      "setp.gt.and.s32 dcool, %3, %13, dh;\n\t"
      "setp.gt.and.s32 dapp, %1, %8, dgo;\n\t"      // scale.go:89
      "@dgo mad.lo.s32 %3, %11, %14, 0;\n\t"        // main.go:76   ld = i
      "@dcool mad.lo.s32 %0, %14, 32, %0;\n\t"
      "@dgo mad.lo.s32 %0, %14, 8, %0;\n\t"
      "@dapp mad.lo.s32 %0, %14, -6, %0;\n\t"
      "sub.s32 u, %1, %10;\n\t"                     // scale.go:94
      "max.s32 u, u, %8;\n\t"                       // scale.go:95-97
      "@dapp mad.lo.s32 %1, u, %14, 0;\n\t"
      "}"
      : "=r"(ev), "+r"(s.r), "+r"(s.lu), "+r"(s.ld)
      : "l"(n), "l"(c.up_msgs), "l"(c.down_msgs), "r"(c.max32), "r"(c.min32), "r"(c.up32),
        "r"(c.down32), "r"(i), "r"(iu), "r"(id), "r"(one));
  return ev;
}

// Trajectory fingerprint step (ksa_summary.hash, include/ksa_replay.h):
//   h <- h * M + ((uint64)event << 32 | (uint32)replicas)
// Three IMADs on the FMA pipe; the addend needs no shifting or masking.
#define KSA_HASH_MUL 0x9E3779B97F4A7C15ull
__device__ __forceinline__ unsigned long long ksa_hash_step(unsigned long long h, int r, unsigned ev) {
  return h * KSA_HASH_MUL + (((unsigned long long)ev << 32) | (unsigned long long)(unsigned)r);
}
