// ksa_sweep.cu — ksa_replay_sweep: the same N traces replayed under P policies (the parameter grid
// of main.go:83-91), one ksa_summary per (policy, trace).
//
// Thread = one trace x a register tile of R policies: each depth value is loaded once and fed to
// R independent state machines (ksa_policy.cuh), so the depth block is re-read P/R times in total
// and stays resident in the 126 MB L2. Unlike the single-policy replay this kernel is bound by
// integer issue, not HBM: compulsory bytes per decision are ~0.01. The state
// machine therefore uses ksa_step_fma / ksa_step_fma_narrow (conditional moves as IMADs on the FMA
// pipe, as few compares as the policy grid allows) and the packed event counters of
// ksa_summary_tick. Policy tiles are split by kind (narrow or not) into two launches, and the grid
// walks L2-sized groups of traces through all policy tiles (see the kernel).
#include "ksa_internal.h"

#include <cstdlib>
#include <algorithm>
#include <mutex>

namespace {

struct SweepArgs {
  const ksa_policy_k* policies;  // [P] device
  const long long* depth;
  long long ld;
  long long N;
  int P, T, A;
  int one;  // always 1 (see ksa_step_fma)
  int split;  // register-tile kernels: 1 = run only the policy tiles whose kind (narrow or not) is this kernel's
  const int* replicas0;
  ksa_summary* summary;  // [P][N]
};

// Whether ksa_step_fma_narrow is exact for a policy: thresholds ordered and inside the range where
// comparing the saturated sum is the same as comparing the sum. Host and device use the same test.
__host__ __device__ inline bool ksa_policy_is_narrow(long long up, long long down) {
  return up > down && up > (long long)INT_MIN && up <= (long long)INT_MAX && down >= (long long)INT_MIN &&
         down < (long long)INT_MAX;
}

// Kind of a policy tile (R consecutive policies, the last tile padded with copies of policy P-1).
// Host (sweep_tile_kinds) and device (the split test in the kernel) use this one function.
enum { TILE_GENERAL = 0, TILE_NARROW = 1, TILE_SHARED = 2 };
template <int R>
__host__ __device__ inline int ksa_tile_kind(const long long (&up)[R], const long long (&down)[R], bool allow_shared) {
  bool narrow = true, shared = allow_shared && up[0] > down[0];
  for (int q = 0; q < R; ++q) {
    narrow = narrow && ksa_policy_is_narrow(up[q], down[q]);
    shared = shared && up[q] == up[0] && down[q] == down[0];
  }
  return shared ? TILE_SHARED : (narrow ? TILE_NARROW : TILE_GENERAL);
}

// KIND = the kind of tile the kernel is compiled for:
//  TILE_NARROW  every policy of the tile is narrow: the step is ksa_step_fma_narrow on the attribute
//               sum saturated once per tick, 6 compares per decision instead of 9.
//  TILE_SHARED  all policies of the tile have the same two thresholds, scale_up_messages >
//               scale_down_messages (the usual case: grids vary the thresholds slowest): the two
//               threshold tests are made once per (trace, tick), on the 64-bit sum, and shared by
//               the R machines (ksa_step_fma_shared, 4 compares per decision).
// A grid that mixes kinds is launched once per kind present with a.split != 0 (1: the three kinds,
// 2: shared tiles not recognised, KSA_SWEEP_NOSHARED); each block then returns at once unless its
// tile is of the kernel's kind.
template <int A, int R, int D, int BLOCK, int MINB, int CHAIN, int KIND>
__global__ void __launch_bounds__(BLOCK, MINB) ksa_sweep_kernel(const __grid_constant__ SweepArgs a) {
  constexpr bool FAST = KIND == TILE_NARROW;
  constexpr bool SHARED = KIND == TILE_SHARED;
  // Block mapping: grid = (trace blocks of a group, policy tiles, groups). Blocks are dispatched x
  // fastest, then y, then z, so a group of gridDim.x trace blocks goes through all policy tiles
  // before the next group starts, and the blocks resident at any time share one L2-sized slice of
  // the depth array. (A flat (trace block, policy tile) grid streams the whole depth array once
  // per policy tile, which only stays in L2 for small arrays.) All block-uniform, so the policy
  // records stay in uniform registers.
  const long long n = ((long long)blockIdx.z * gridDim.x + blockIdx.x) * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const int p0 = blockIdx.y * R;
  const int T = a.T, one = a.one;

  // The R policy records of this block (block-uniform addresses; served from L1 / L2).
  ksa_policy_k pk[R];
#pragma unroll
  for (int q = 0; q < R; ++q) pk[q] = a.policies[min(p0 + q, a.P - 1)];
  if (a.split) {
    long long up[R], down[R];
#pragma unroll
    for (int q = 0; q < R; ++q) up[q] = pk[q].up_msgs, down[q] = pk[q].down_msgs;
    if (ksa_tile_kind<R>(up, down, a.split == 1) != KIND) return;
  }

  ksa_trace_regs s[R];
  ksa_summary_acc m[R];
  const int r0 = a.replicas0[n];
#pragma unroll
  for (int q = 0; q < R; ++q) {
    const ksa_step_consts c(pk[q]);
    s[q].r = r0;
    s[q].su = ksa_local_stamp(pk[q].cu, c.cu, 0, 0);  // fresh process: both stamps at tick 0,
    s[q].sd = ksa_local_stamp(pk[q].cd, c.cd, 0, 0);  // main.go:37-38
    ksa_summary_acc_init(m[q], INT_MAX, INT_MIN, 0, 0ull);
    if (p0 + q < a.P) {  // counters live in the output record; start them at zero
      ksa_summary* z = a.summary + (long long)(p0 + q) * a.N + n;
      z->up_applied = z->down_applied = z->up_noop = z->down_noop = z->up_cool = z->down_cool = 0u;
    }
  }
  auto fold = [&]() {
#pragma unroll
    for (int q = 0; q < R; ++q) ksa_summary_fold(m[q]);
  };
  auto flush = [&]() {
#pragma unroll
    for (int q = 0; q < R; ++q)
      if (p0 + q < a.P) ksa_summary_flush(m[q], a.summary + (long long)(p0 + q) * a.N + n);
  };

  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + n;
  auto load_sum = [&](int t) {
    unsigned long long acc = (unsigned long long)__ldg(dp + (long long)t * row);
#pragma unroll
    for (int p = 1; p < A; ++p) acc += (unsigned long long)__ldg(dp + (long long)t * row + p * a.ld);
    return (long long)acc;  // sqs.go:56-64, wrapping
  };
  long long buf[D];
#pragma unroll
  for (int j = 0; j < D; ++j) buf[j] = j < T ? load_sum(j) : 0;

  // One tick of the R machines. `tt` = tick (0-based), j = its ring slot, prefetch = tick tt + D exists.
  auto tick = [&](const int tt, const int j, const bool prefetch, const int phase) {
    const long long msgs = buf[j];
    // The narrow and shared kernels keep (part of) the sum in the ring slot's registers for the
    // whole tick; refilling the slot here would go through a temporary that ptxas copies back at the
    // end of the tick, i.e. it would wait for the load it has just issued. They refill after the
    // machines.
    if (KIND == TILE_GENERAL && prefetch) buf[j] = load_sum(tt + D);
    const int i = tt + 1;
    // CHAIN > 0: the R machines run in groups of CHAIN. The first compare of a group is made
    // to wait for the last event of the previous group through a xor that changes nothing
    // (bit 4 of a wide event code is always 0, which ptxas cannot know). Without it ptxas
    // interleaves all R machines, runs out of its 7 predicate registers and spills them into
    // a general register (13 LOP3 + 3 P2R per decision at R = 4).
    unsigned m_lo = (unsigned)(unsigned long long)msgs;
    const unsigned m_hi = (unsigned)((unsigned long long)msgs >> 32);
    int m32 = FAST ? ksa_saturate_s32(msgs) : 0;
    unsigned ev_last = 0u;
    // SHARED: main.go:51,65 once for the whole tile (see ksa_step_fma_shared)
    int iu = 0, id = 0;
    unsigned e0 = 0u, e0n = 0u;
    if constexpr (SHARED) {
      const bool uh = msgs >= pk[0].up_msgs, dh = msgs <= pk[0].down_msgs;
      iu = uh ? i : INT_MIN;
      id = dh ? i : INT_MIN;
      e0 = uh ? KSA_W_UP_COOL : (dh ? KSA_W_DOWN_COOL : 0u);
      e0n = e0 + (KSA_W_UP_NOOP - KSA_W_UP_COOL);
    }
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const ksa_step_consts c(pk[q]);
      if (CHAIN > 0 && q > 0 && q % CHAIN == 0) {
        if constexpr (SHARED)
          asm volatile("lop3.b32 %0, %0, %1, 16, 0x78;" : "+r"(iu) : "r"(ev_last));  // iu ^ (ev & 16)
        else if constexpr (FAST)
          asm volatile("lop3.b32 %0, %0, %1, 16, 0x78;" : "+r"(m32) : "r"(ev_last));  // m32 ^ (ev & 16)
        else
          asm volatile("lop3.b32 %0, %0, %1, 16, 0x78;" : "+r"(m_lo) : "r"(ev_last));
      }
      unsigned ev;  // wide event code
      if constexpr (SHARED)
        ev = ksa_step_fma_shared(c, iu, id, e0, e0n, i, one, s[q]);
      else if constexpr (FAST)
        ev = ksa_step_fma_narrow((int)pk[q].up_msgs, (int)pk[q].down_msgs, c, m32, i, one, s[q]);
      else
        ev = ksa_step_fma<true>(c, (long long)(((unsigned long long)m_hi << 32) | m_lo), i, one, s[q]);
      ev_last = ev;
      // (pairing the min / max updates of two ticks, as the replay kernels do, costs a register
      // per machine: with the narrow step it measured slower, 74.0 vs 71.4 ms; the shared step
      // holds fewer values per machine and takes it)
      ksa_summary_tick<true>(m[q], s[q].r, ev, one, phase);
    }
    if (KIND != TILE_GENERAL && prefetch) buf[j] = load_sum(tt + D);
  };
  int pending = 0, folds = 0;
  auto count_folds = [&]() {
    pending += D;
    if (pending > KSA_FOLD_TICKS - D) {
      pending = 0;
      if (++folds == KSA_FLUSH_FOLDS) {
        flush();
        folds = 0;
      } else {
        fold();
      }
    }
  };
  int t = 0;
  // Steady state: every tick of the group exists and so does the tick D ahead of it (no guards).
#pragma unroll 1
  for (; t + 2 * D <= T; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      tick(t + j, j, true, (SHARED && D % 2 == 0) ? (j % 2 ? KSA_TICK_SECOND : KSA_TICK_FIRST) : KSA_TICK_SINGLE);
    count_folds();
  }
  // Drain: at most 2D-1 ticks left; guard both the tick and its prefetch.
#pragma unroll 1
  for (; t < T; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j)
      if (t + j < T) tick(t + j, j, t + j + D < T, KSA_TICK_SINGLE);
    count_folds();
  }
  flush();
#pragma unroll
  for (int q = 0; q < R; ++q) {
    if (p0 + q < a.P) {
      ksa_summary* z = a.summary + (long long)(p0 + q) * a.N + n;
      if (m[q].bad) {  // a replica count left +-KSA_SUM_SAFE: redo the sum exactly (ksa_summary_acc)
        const ksa_step_consts c(pk[q]);
        ksa_trace_regs s0;
        s0.r = r0;
        s0.su = ksa_local_stamp(pk[q].cu, c.cu, 0, 0);
        s0.sd = ksa_local_stamp(pk[q].cd, c.cd, 0, 0);
        m[q].sum = ksa_exact_sum(pk[q], dp, row, a.ld, A, nullptr, 0, T, s0);
      }
      z->final_replicas = s[q].r;
      z->min_replicas = m[q].rmin;
      z->max_replicas = m[q].rmax;
      z->poll_failed = z->up_failed = z->down_failed = 0u;
      z->sum_replicas = m[q].sum;
      z->hash = ((unsigned long long)m[q].hb << 32) | m[q].ha;
    }
  }
}

// Second layout: lane = policy. A block replays V adjacent traces under BLOCK consecutive policies;
// the depth value of a (trace, tick) is block-uniform (one broadcast load), the policy constants
// and the whole state of a (policy, trace) pair are per-lane registers. One or two state machines
// per thread, so ptxas never has more than a few predicates live (the register-tile kernel above
// runs four machines per thread and, in SASS, spends 13 LOP3 + 3 P2R per decision spilling
// predicates into a general register; profiles/r01_ncu_sweep.txt: 48.5 instructions per decision).
// Latency is hidden by the ~40 resident warps per SM instead of by ILP inside a thread.
template <int A, int V, int D, int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) ksa_sweep_lane_kernel(const __grid_constant__ SweepArgs a) {
  const int pblocks = (a.P + BLOCK - 1) / BLOCK;
  const int pb = (int)(blockIdx.x % (unsigned)pblocks);        // policy block: fastest, so that
  const long long n0 = (long long)(blockIdx.x / (unsigned)pblocks) * V;  // co-running blocks share traces
  const int p = pb * BLOCK + threadIdx.x;
  const bool live = p < a.P;
  const int T = a.T, one = a.one;
  const ksa_policy_k pk = a.policies[live ? p : a.P - 1];
  const ksa_step_consts c(pk);

  ksa_trace_regs s[V];
  ksa_summary_acc m[V];
  bool have[V];
#pragma unroll
  for (int v = 0; v < V; ++v) {
    have[v] = live && (n0 + v < a.N);
    s[v].r = a.replicas0[min(n0 + v, a.N - 1)];
    s[v].su = ksa_local_stamp(pk.cu, c.cu, 0, 0);  // fresh process: both stamps at tick 0,
    s[v].sd = ksa_local_stamp(pk.cd, c.cd, 0, 0);  // main.go:37-38
    ksa_summary_acc_init(m[v], INT_MAX, INT_MIN, 0, 0ull);
    if (have[v]) {  // counters live in the output record; start them at zero
      ksa_summary* z = a.summary + (long long)p * a.N + n0 + v;
      z->up_applied = z->down_applied = z->up_noop = z->down_noop = z->up_cool = z->down_cool = 0u;
    }
  }
  auto fold = [&]() {
#pragma unroll
    for (int v = 0; v < V; ++v) ksa_summary_fold(m[v]);
  };
  auto flush = [&]() {
#pragma unroll
    for (int v = 0; v < V; ++v)
      if (have[v]) ksa_summary_flush(m[v], a.summary + (long long)p * a.N + n0 + v);
  };

  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + n0;  // block-uniform
  auto load_sum = [&](int t, int v) {
    const long long* q = dp + (long long)t * row + (n0 + v < a.N ? v : 0);
    unsigned long long acc = (unsigned long long)__ldg(q);
#pragma unroll
    for (int k = 1; k < A; ++k) acc += (unsigned long long)__ldg(q + k * a.ld);
    return (long long)acc;  // sqs.go:56-64, wrapping
  };
  long long buf[D][V];
#pragma unroll
  for (int j = 0; j < D; ++j)
#pragma unroll
    for (int v = 0; v < V; ++v) buf[j][v] = j < T ? load_sum(j, v) : 0;

  int pending = 0, folds = 0;
#pragma unroll 1
  for (int t = 0; t < T; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const int tt = t + j;
      if (tt < T) {
        long long msgs[V];
#pragma unroll
        for (int v = 0; v < V; ++v) msgs[v] = buf[j][v];
        if (tt + D < T) {
#pragma unroll
          for (int v = 0; v < V; ++v) buf[j][v] = load_sum(tt + D, v);
        }
        const int i = tt + 1;
#pragma unroll
        for (int v = 0; v < V; ++v) {
          const unsigned ev = ksa_step_fma<true>(c, msgs[v], i, one, s[v]);  // wide event code
          ksa_summary_tick<true>(m[v], s[v].r, ev, one);
        }
      }
    }
    pending += D;
    if (pending > KSA_FOLD_TICKS - D) {
      pending = 0;
      if (++folds == KSA_FLUSH_FOLDS) {
        flush();
        folds = 0;
      } else {
        fold();
      }
    }
  }
  flush();

#pragma unroll
  for (int v = 0; v < V; ++v) {
    if (have[v]) {
      ksa_summary* z = a.summary + (long long)p * a.N + n0 + v;
      if (m[v].bad) {  // see ksa_sweep_kernel
        ksa_trace_regs s0;
        s0.r = a.replicas0[n0 + v];
        s0.su = ksa_local_stamp(pk.cu, c.cu, 0, 0);
        s0.sd = ksa_local_stamp(pk.cd, c.cd, 0, 0);
        m[v].sum = ksa_exact_sum(pk, dp + v, row, a.ld, A, nullptr, 0, T, s0);
      }
      z->final_replicas = s[v].r;
      z->min_replicas = m[v].rmin;
      z->max_replicas = m[v].rmax;
      z->poll_failed = z->up_failed = z->down_failed = 0u;
      z->sum_replicas = m[v].sum;
      z->hash = ((unsigned long long)m[v].hb << 32) | m[v].ha;
    }
  }
}

typedef void (*sweep_fn)(const SweepArgs);
struct SweepVariant {
  int id;     // stable across builds (KSA_SWEEP_VARIANT=<id>)
  int R;      // register-tile kernel: policies per thread; lane kernel: 0
  int V;      // lane kernel: traces per thread
  int block;
  const char* name;
  sweep_fn fn[3][3];  // [tile kind][A-1]; the lane kernels have one form for every kind
};
#define KSA_SWEEP_CH(ID, R, D, BLOCK, MINB, CH)                                                    \
  {ID, R, 0, BLOCK, "R" #R "_D" #D "_B" #BLOCK "x" #MINB "_C" #CH,                                 \
   {{ksa_sweep_kernel<1, R, D, BLOCK, MINB, CH, TILE_GENERAL>, ksa_sweep_kernel<2, R, D, BLOCK, MINB, CH, TILE_GENERAL>, \
     ksa_sweep_kernel<3, R, D, BLOCK, MINB, CH, TILE_GENERAL>},                                    \
    {ksa_sweep_kernel<1, R, D, BLOCK, MINB, CH, TILE_NARROW>, ksa_sweep_kernel<2, R, D, BLOCK, MINB, CH, TILE_NARROW>, \
     ksa_sweep_kernel<3, R, D, BLOCK, MINB, CH, TILE_NARROW>},                                     \
    {ksa_sweep_kernel<1, R, D, BLOCK, MINB, CH, TILE_SHARED>, ksa_sweep_kernel<2, R, D, BLOCK, MINB, CH, TILE_SHARED>, \
     ksa_sweep_kernel<3, R, D, BLOCK, MINB, CH, TILE_SHARED>}}}
#define KSA_SWEEP(ID, R, D, BLOCK, MINB) KSA_SWEEP_CH(ID, R, D, BLOCK, MINB, 0)
#define KSA_LANE3(V, D, BLOCK, MINB)                                                             \
  {ksa_sweep_lane_kernel<1, V, D, BLOCK, MINB>, ksa_sweep_lane_kernel<2, V, D, BLOCK, MINB>,     \
   ksa_sweep_lane_kernel<3, V, D, BLOCK, MINB>}
#define KSA_SWEEP_LANE(ID, V, D, BLOCK, MINB)                                                    \
  {ID, 0, V, BLOCK, "LANE_V" #V "_D" #D "_B" #BLOCK "x" #MINB,                                   \
   {KSA_LANE3(V, D, BLOCK, MINB), KSA_LANE3(V, D, BLOCK, MINB), KSA_LANE3(V, D, BLOCK, MINB)}}
// Launch shapes. The product build has the defaults only; -DKSA_TUNING_VARIANTS adds the points that
// were measured (profiles/r01_config4_sweep_tune.txt). KSA_SWEEP_VARIANT=<id> (environment)
// overrides the default for measurements and tests.
const SweepVariant kSweepVariants[] = {
    KSA_SWEEP(26, 4, 3, 128, 4),  // A=1: every kind
    KSA_SWEEP(33, 4, 1, 256, 2),  // A=2,3: every kind (three 2-KB rows per tick want the wider block)
    KSA_SWEEP_LANE(9, 2, 4, 128, 8),  // the second layout (lane = policy), kept as a cross-check
#ifdef KSA_TUNING_VARIANTS
    KSA_SWEEP(0, 4, 4, 128, 3), KSA_SWEEP(1, 2, 4, 128, 6), KSA_SWEEP(2, 4, 2, 128, 4), KSA_SWEEP(3, 2, 4, 256, 3),
    KSA_SWEEP(4, 1, 8, 128, 8), KSA_SWEEP(5, 8, 2, 128, 2), KSA_SWEEP(6, 4, 4, 64, 6), KSA_SWEEP(7, 3, 4, 128, 4),
    KSA_SWEEP_LANE(8, 1, 4, 128, 10), KSA_SWEEP_LANE(10, 1, 8, 128, 8),
    KSA_SWEEP_LANE(11, 2, 2, 128, 8), KSA_SWEEP_LANE(12, 1, 4, 256, 4), KSA_SWEEP_LANE(13, 2, 4, 256, 4),
    KSA_SWEEP_LANE(14, 1, 2, 128, 12), KSA_SWEEP_LANE(15, 1, 4, 64, 16),
    // 16..: register tile with chained groups
    KSA_SWEEP_CH(16, 4, 4, 128, 3, 2), KSA_SWEEP_CH(17, 4, 2, 128, 4, 2), KSA_SWEEP_CH(18, 4, 4, 128, 4, 2),
    KSA_SWEEP_CH(19, 4, 4, 128, 3, 1), KSA_SWEEP_CH(20, 8, 2, 128, 2, 2), KSA_SWEEP_CH(21, 8, 2, 128, 2, 4),
    KSA_SWEEP_CH(22, 6, 2, 128, 3, 2), KSA_SWEEP_CH(23, 4, 4, 64, 6, 2),
    // 24..: more unchained R=4 points (the narrow step needs fewer predicates)
    KSA_SWEEP(24, 4, 1, 128, 4), KSA_SWEEP(25, 4, 2, 64, 8), KSA_SWEEP(27, 4, 2, 256, 2),
    KSA_SWEEP(28, 4, 2, 128, 5), KSA_SWEEP(29, 4, 2, 128, 3),
    // 30..: 256-thread blocks (2 KB depth rows per block and tick; better with three attributes)
    KSA_SWEEP_CH(30, 4, 2, 256, 2, 2), KSA_SWEEP_CH(31, 4, 4, 256, 2, 2), KSA_SWEEP(32, 4, 3, 256, 2),
    KSA_SWEEP(34, 4, 1, 512, 1), KSA_SWEEP(35, 4, 2, 512, 1), KSA_SWEEP_CH(36, 4, 1, 256, 2, 2), KSA_SWEEP(37, 4, 1, 128, 5),
    // 38..: shapes for the shared-threshold tiles (fewer live predicates: deeper rings / more blocks fit)
    KSA_SWEEP(38, 4, 2, 128, 6), KSA_SWEEP(39, 4, 4, 128, 4), KSA_SWEEP(40, 8, 2, 128, 2), KSA_SWEEP(41, 8, 1, 128, 3),
    KSA_SWEEP(42, 6, 2, 128, 3), KSA_SWEEP(43, 4, 2, 64, 10), KSA_SWEEP(44, 8, 2, 64, 5),
#endif
};
const SweepVariant* find_sweep_variant(int id) {
  for (const SweepVariant& v : kSweepVariants)
    if (v.id == id) return &v;
  return nullptr;
}
// Defaults per tile kind and number of attributes (index A-1). All have R = 4, which the split launch
// of a mixed grid needs (tiles of the kernels must coincide).
//   A=1: R4 D3 128x4 for every kind: 58.4 ms on configs[3] (profiles/r02_sweep_tune.txt; R4 D2 128x4,
//        the round-1 choice for narrow tiles: 61.6 ms; R4 D1 128x5 for general tiles: 60.3 ms)
//   A=2,3: R4 D1 256x2 (three 2-KB rows per tick want the wider block; profiles/r01_config4_sweep_tune.txt)
constexpr int kDefaultSweepId[3][3] = {{26, 33, 33}, {26, 33, 33}, {26, 33, 33}};

// Grid of a sweep launch, or false if it does not fit one launch. Register-tile kernels:
// (gw trace blocks, policy tiles, groups of gw trace blocks), see the block mapping in the kernel;
// gw is chosen so that a group's slice of the depth array is about 48 MB.
bool sweep_grid(const SweepVariant& v, int P, long long N, int A, int T, dim3* grid) {
  if (v.R > 0) {
    const long long gy = (P + v.R - 1) / v.R;
    if (gy > 65535) return false;
    const long long gx = (N + v.block - 1) / v.block;
    long long gw = 48000000ll / ((long long)v.block * 8 * A * T);
    if (const char* e = std::getenv("KSA_SWEEP_GW"))  // tuning knob: trace blocks per group
      if (std::atoi(e) > 0) gw = std::atoi(e);
    gw = std::max<long long>(1, std::min(gw, gx));
    const long long gz = (gx + gw - 1) / gw;
    if (gz > 65535) return false;
    *grid = dim3((unsigned)gw, (unsigned)gy, (unsigned)gz);
  } else {
    const long long blocks = (long long)((P + v.block - 1) / v.block) * ((N + v.V - 1) / v.V);
    if (blocks > 0x7fffffffll) return false;
    *grid = dim3((unsigned)blocks);
  }
  return true;
}

// Which kinds of policy tiles (ksa_tile_kind; R consecutive policies, the last tile padded with
// copies of policy P-1 as in the kernel) a grid has: bit k set = some tile is of kind k.
template <int R>
unsigned sweep_tile_kinds_r(const ksa_policy* policies, int P, bool allow_shared) {
  unsigned kinds = 0u;
  for (int p0 = 0; p0 < P; p0 += R) {
    long long up[R], down[R];
    for (int q = 0; q < R; ++q) {
      const ksa_policy& pol = policies[std::min(p0 + q, P - 1)];
      up[q] = pol.scale_up_messages, down[q] = pol.scale_down_messages;
    }
    kinds |= 1u << ksa_tile_kind<R>(up, down, allow_shared);
  }
  return kinds;
}
unsigned sweep_tile_kinds(const ksa_policy* policies, int P, int R, bool allow_shared) {
  switch (R) {
    case 1: return sweep_tile_kinds_r<1>(policies, P, allow_shared);
    case 2: return sweep_tile_kinds_r<2>(policies, P, allow_shared);
    case 3: return sweep_tile_kinds_r<3>(policies, P, allow_shared);
    case 4: return sweep_tile_kinds_r<4>(policies, P, allow_shared);
    case 6: return sweep_tile_kinds_r<6>(policies, P, allow_shared);
    case 8: return sweep_tile_kinds_r<8>(policies, P, allow_shared);
    default: return 1u << TILE_GENERAL;
  }
}

// Launch the sweep: one launch per kind of tile the grid has, each kernel on the tiles of its kind.
// KSA_SWEEP_GENERAL=1 (environment) forces the general kernels for every tile, KSA_SWEEP_NOSHARED=1
// classifies without the shared kind (shared tiles then run as narrow or general ones), for
// measurements and tests.
int sweep_launch(const ksa_policy* policies, SweepArgs a, int A, cudaStream_t st, const char* what) {
  const SweepVariant* v[3];
  for (int k = 0; k < 3; ++k) v[k] = find_sweep_variant(kDefaultSweepId[k][A - 1]);
  if (const char* e = std::getenv("KSA_SWEEP_VARIANT")) {
    const SweepVariant* want = find_sweep_variant(std::atoi(e));
    if (!want) return ksa_fail(KSA_EINVAL, "unknown sweep variant %s (KSA_SWEEP_VARIANT)", e);
    v[0] = v[1] = v[2] = want;
  }
  if (const char* e = std::getenv("KSA_SWEEP_SHARED_VARIANT")) {  // tuning knob: shape of the shared-tile kernel only
    const SweepVariant* want = find_sweep_variant(std::atoi(e));
    if (!want || want->R != v[1]->R) return ksa_fail(KSA_EINVAL, "bad KSA_SWEEP_SHARED_VARIANT %s", e);
    v[2] = want;
  }
  int force = 0;
  if (const char* e = std::getenv("KSA_SWEEP_GENERAL")) force = std::atoi(e);
  // Tiles of the kernels must coincide for the split; the lane kernels have one form.
  unsigned kinds = 1u << TILE_GENERAL;
  bool allow_shared = true;
  if (const char* e = std::getenv("KSA_SWEEP_NOSHARED")) allow_shared = std::atoi(e) == 0;
  if (v[0]->R > 0 && v[0]->R == v[1]->R && v[0]->R == v[2]->R && force != 1)
    kinds = sweep_tile_kinds(policies, a.P, v[0]->R, allow_shared);
  // a kernel tests the kind of its tile only when another kernel has tiles too
  a.split = (kinds & (kinds - 1)) != 0 ? (allow_shared ? 1 : 2) : 0;
  dim3 grid;
  for (int k = 0; k < 3; ++k) {
    if (!(kinds & (1u << k))) continue;
    if (!sweep_grid(*v[k], a.P, a.N, A, a.T, &grid)) return ksa_fail(KSA_ELIMIT, "P x N too large for one launch");
    v[k]->fn[k][A - 1]<<<grid, v[k]->block, 0, st>>>(a);
    KSA_LAUNCH_CHECK(what);
  }
  return KSA_OK;
}

// The kernel-side policy table is written by small kernels that carry the records in their
// parameter block (640 records = 30 KB per launch; sm_100a takes 32 KB of parameters), not by a
// host-to-device copy: a copy from a host buffer would still be reading that buffer when the call
// has returned if the stream is being captured into a graph (the copy becomes a node that runs at
// graph launch), while kernel parameters are taken by value at the call.
constexpr int kPolicySlab = 640;
struct PolicySlab {
  ksa_policy_k rec[kPolicySlab];
};
__global__ void __launch_bounds__(128) policy_fill_kernel(const __grid_constant__ PolicySlab slab,
                                                          ksa_policy_k* __restrict__ dst, int count) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i < count) dst[i] = slab.rec[i];
}

// Validate and convert policies [0, P) and write them to table_dev[0, P) on `st`. No allocation.
int upload_policies(const ksa_policy* policies, int P, int A, ksa_policy_k* table_dev, cudaStream_t st) {
  PolicySlab slab;  // 30 KB of stack
  for (int base = 0; base < P; base += kPolicySlab) {
    const int count = std::min(kPolicySlab, P - base);
    for (int q = 0; q < count; ++q) {
      int rc = ksa_policy_to_k(&policies[base + q], A, &slab.rec[q]);
      if (rc != KSA_OK) return rc;
      if (policies[base + q].attr_mask != policies[0].attr_mask)
        return ksa_fail(KSA_EINVAL, "policy %d: attr_mask differs from policy 0", base + q);
    }
    policy_fill_kernel<<<(count + 127) / 128, 128, 0, st>>>(slab, table_dev + base, count);
    KSA_LAUNCH_CHECK("policy_fill_kernel");
  }
  return KSA_OK;
}

}  // namespace

// For ksa_variant_name(4, id): name of a sweep launch shape, or NULL if this build does not have it.
const char* ksa_sweep_variant_name(int id) {
  if (id == -1) id = kDefaultSweepId[TILE_SHARED][0];
  const SweepVariant* v = find_sweep_variant(id);
  return v ? v->name : nullptr;
}

extern "C" int ksa_replay_sweep(const ksa_policy* policies, int32_t P, void* policies_dev,
                                const int64_t* depth_dev, int64_t T, int32_t A, int64_t N, int64_t ld,
                                const int32_t* replicas0_dev, ksa_summary* summary_dev, void* stream) {
  if (!policies || !policies_dev || !depth_dev || !replicas0_dev || !summary_dev)
    return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (P < 1 || P > (1 << 20)) return ksa_fail(KSA_EINVAL, "P=%d outside [1, 2^20]", P);
  if (T < 1 || T > KSA_MAX_CHUNK_TICKS) return ksa_fail(KSA_EINVAL, "T=%lld outside [1, 2^30 - 64]", (long long)T);
  if (N < 1 || N > 0x7fffffffll || ld < N) return ksa_fail(KSA_EINVAL, "bad N=%lld / ld=%lld", (long long)N, (long long)ld);
  if (A < 1 || A > KSA_MAX_ATTRS) return ksa_fail(KSA_EINVAL, "A=%d outside [1,3]", A);
  cudaStream_t st = (cudaStream_t)stream;
  int rc = upload_policies(policies, P, A, (ksa_policy_k*)policies_dev, st);
  if (rc != KSA_OK) return rc;
  SweepArgs a;
  a.policies = (const ksa_policy_k*)policies_dev;
  a.depth = (const long long*)depth_dev;
  a.ld = ld;
  a.N = N;
  a.P = P;
  a.T = (int)T;
  a.A = A;
  a.one = 1;
  a.split = 0;
  a.replicas0 = replicas0_dev;
  a.summary = summary_dev;
  return sweep_launch(policies, a, A, st, "ksa_sweep_kernel");
}

// ---- ksa_replay_sweep_static: the sweep for callers that must not allocate -----------------------
//
// Same kernel, same results as ksa_replay_sweep; the only difference is where the kernel-side
// policy table (the sweep's scratch memory) lives. Here it is a fixed array inside the library's
// device image, KSA_SWEEP_STATIC_MAX_POLICIES records long, so the call needs no buffer from the
// caller and performs no device or host allocation of its own. It is meant for hosts that have no
// safe way to get 48*P bytes of device memory at call time: the GPU's memory already belongs to
// another framework's pool, the host is Go over cgo and owns no CUDA allocator, or the sweep runs
// in a loop that must not be stalled by cudaMalloc / cudaFree (include/ksa_replay.h has the
// contract: size limit, first-call set-up, shared table).

namespace {

// 4096 x 48 B = 192 KB of device memory on every GPU that loads the library, held until exit.
static __device__ __align__(256) unsigned char g_static_policy_table[KSA_SWEEP_STATIC_MAX_POLICIES * sizeof(ksa_policy_k)];

constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mutex;                          // guards the array below
void* g_static_table_addr[kMaxStaticDevices] = {};  // device address of the table, per device

}  // namespace

extern "C" int ksa_replay_sweep_static(const ksa_policy* policies, int32_t P,
                                       const int64_t* depth_dev, int64_t T, int32_t A, int64_t N,
                                       int64_t ld, const int32_t* replicas0_dev,
                                       ksa_summary* summary_dev, void* stream) {
  if (!policies || !depth_dev || !replicas0_dev || !summary_dev)
    return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (P < 1) return ksa_fail(KSA_EINVAL, "P=%d < 1", P);
  if (P > KSA_SWEEP_STATIC_MAX_POLICIES)
    return ksa_fail(KSA_ELIMIT,
                    "P=%d exceeds the built-in policy table of ksa_replay_sweep_static (%d policies); "
                    "split the grid or use ksa_replay_sweep with a caller-provided table",
                    P, KSA_SWEEP_STATIC_MAX_POLICIES);
  if (T < 1 || T > KSA_MAX_CHUNK_TICKS) return ksa_fail(KSA_EINVAL, "T=%lld outside [1, 2^30 - 64]", (long long)T);
  if (N < 1 || N > 0x7fffffffll || ld < N) return ksa_fail(KSA_EINVAL, "bad N=%lld / ld=%lld", (long long)N, (long long)ld);

  int device = -1;
  KSA_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= kMaxStaticDevices)
    return ksa_fail(KSA_ELIMIT, "device ordinal %d outside [0, %d)", device, kMaxStaticDevices);

  if (A < 1 || A > KSA_MAX_ATTRS) return ksa_fail(KSA_EINVAL, "A=%d outside [1,3]", A);
  // Set-up, once per device, with that device current: in host code the array's name is only a
  // handle for the runtime, the address the kernel must be given comes from the runtime.
  void* table = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    table = g_static_table_addr[device];
    if (!table) {
      KSA_CUDA(cudaGetSymbolAddress(&table, g_static_policy_table));
      g_static_table_addr[device] = table;
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  // Records [0, P) are rewritten by every call (by kernels on `st`, ahead of the sweep kernel) and
  // nothing past P-1 is read, so nothing left over from an earlier call needs clearing. The
  // conversion buffer is on this thread's stack: no host state is shared between calls.
  int rc = upload_policies(policies, P, A, (ksa_policy_k*)table, st);
  if (rc != KSA_OK) return rc;
  SweepArgs a;
  a.policies = (const ksa_policy_k*)table;
  a.depth = (const long long*)depth_dev;
  a.ld = ld;
  a.N = N;
  a.P = P;
  a.T = (int)T;
  a.A = A;
  a.one = 1;
  a.split = 0;
  a.replicas0 = replicas0_dev;
  a.summary = summary_dev;
  return sweep_launch(policies, a, A, st, "ksa_sweep_kernel (static table)");
}

// ---- ksa_sweep_reduce: score a sweep and pick policies (extension; the reference has no cost) -----
//
// One pass over the [P][N] summaries. A block owns 32 adjacent traces (lane = trace) and its
// warps split the policies (warp w takes p = w, w + kReduceWarps, ...), so that even a 4096-trace
// grid fills the GPU: N/32 blocks of 32 warps, each warp keeping kReduceUnroll 2-KB rows (32
// records of 64 B) in flight. Per policy the warp's 32 costs are summed with shuffles and added to
// policy_cost[p] with one 64-bit atomic; wrapping int64 addition is associative, so the totals are
// exact whatever the order of the atomics. Per trace every warp keeps the argmin over its own
// policies (ascending p, strict <), and warp 0 merges the candidates of all warps through shared memory,
// cost first, then the lower p.

namespace {

template <int kReduceWarps, int kReduceUnroll>
__global__ void __launch_bounds__(kReduceWarps * 32) sweep_reduce_kernel(
    const ksa_summary* __restrict__ summary, int P, long long N, const __grid_constant__ ksa_cost w,
    long long* __restrict__ policy_cost, int* __restrict__ best_policy, long long* __restrict__ best_cost) {
  __shared__ long long s_best[kReduceWarps][32];
  __shared__ int s_arg[kReduceWarps][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long n = (long long)blockIdx.x * 32 + lane;
  const bool live = n < N;
  long long best = 0;
  int arg = -1;
  for (int p0 = warp; p0 < P; p0 += kReduceWarps * kReduceUnroll) {
    // the record as four 16-byte loads (it is 64-byte aligned); all rows of the round first
    uint4 a[kReduceUnroll], b[kReduceUnroll], d[kReduceUnroll], e[kReduceUnroll];
#pragma unroll
    for (int q = 0; q < kReduceUnroll; ++q) {
      const int p = p0 + q * kReduceWarps;
      a[q] = b[q] = d[q] = e[q] = make_uint4(0, 0, 0, 0);
      if (live && p < P) {
        const uint4* rec = reinterpret_cast<const uint4*>(summary + (long long)p * N + n);
        a[q] = __ldg(rec), b[q] = __ldg(rec + 1), d[q] = __ldg(rec + 2), e[q] = __ldg(rec + 3);
      }
    }
#pragma unroll
    for (int q = 0; q < kReduceUnroll; ++q) {
      const int p = p0 + q * kReduceWarps;
      // fields by byte offset: 12 up_applied, 16 down_applied, 20 up_noop, 28 up_cool,
      // 32 down_cool, 48 sum_replicas
      const unsigned up_applied = a[q].w, down_applied = b[q].x, up_noop = b[q].y, up_cool = b[q].w,
                     down_cool = d[q].x;
      const unsigned long long sum_replicas = (unsigned long long)e[q].x | ((unsigned long long)e[q].y << 32);
      const unsigned long long c =
          (unsigned long long)w.w_pod_ticks * sum_replicas +
          (unsigned long long)w.w_saturated * (unsigned long long)up_noop +
          (unsigned long long)w.w_churn * ((unsigned long long)up_applied + (unsigned long long)down_applied) +
          (unsigned long long)w.w_blocked * ((unsigned long long)up_cool + (unsigned long long)down_cool);
      if (live && p < P && (arg < 0 || (long long)c < best)) {
        best = (long long)c;
        arg = p;
      }
      if (policy_cost && p < P) {  // p is warp-uniform; dead lanes carry c == 0
        unsigned long long s = c;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
        if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(policy_cost) + p, s);
      }
    }
  }
  if (!best_policy && !best_cost) return;
  s_best[warp][lane] = best;
  s_arg[warp][lane] = arg;
  __syncthreads();
  if (warp == 0 && live) {
#pragma unroll
    for (int k = 1; k < kReduceWarps; ++k) {
      const long long ob = s_best[k][lane];
      const int oa = s_arg[k][lane];
      if (oa >= 0 && (arg < 0 || ob < best || (ob == best && oa < arg))) {
        best = ob;
        arg = oa;
      }
    }
    if (best_policy) best_policy[n] = arg;
    if (best_cost) best_cost[n] = best;
  }
}

}  // namespace

extern "C" int ksa_sweep_reduce(const ksa_summary* summary_dev, int32_t P, int64_t N, const ksa_cost* cost,
                                int64_t* policy_cost_dev, int32_t* best_policy_dev, int64_t* best_cost_dev,
                                void* stream) {
  if (!summary_dev || !cost) return ksa_fail(KSA_EINVAL, "NULL summary_dev or cost");
  if (P < 1 || N < 1) return ksa_fail(KSA_EINVAL, "P=%d, N=%lld: both must be >= 1", P, (long long)N);
  const long long blocks = (N + 31) / 32;
  if (blocks > 0x7fffffffll) return ksa_fail(KSA_ELIMIT, "N too large for one launch");
  int shape = 0;
  if (const char* e = std::getenv("KSA_REDUCE_SHAPE")) shape = std::atoi(e);  // tuning knob (tools/)
#define KSA_RED(W, U)                                                                                  \
  sweep_reduce_kernel<W, U><<<(unsigned)blocks, W * 32, 0, (cudaStream_t)stream>>>(                       \
      summary_dev, P, N, *cost, (long long*)policy_cost_dev, best_policy_dev, (long long*)best_cost_dev)
  switch (shape) {
#ifdef KSA_TUNING_VARIANTS
    case 1: KSA_RED(16, 8); break;
    case 2: KSA_RED(32, 4); break;
    case 3: KSA_RED(32, 8); break;
    case 4: KSA_RED(8, 8); break;
    case 5: KSA_RED(16, 4); break;
    case 6: KSA_RED(32, 1); break;
    case 7: KSA_RED(32, 3); break;
#endif
    default: KSA_RED(32, 2); break;  // 0.184 ms on 4096 x 4096 records; (16, 4): 0.244 ms (profiles/r01_config4_reduce.txt)
  }
#undef KSA_RED
  KSA_LAUNCH_CHECK("sweep_reduce_kernel");
  return KSA_OK;
}
