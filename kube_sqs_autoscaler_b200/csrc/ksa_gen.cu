// ksa_gen.cu — everything around the hot path that also lives on the device: the synthetic depth
// generator (include/ksa_tracegen.h compiled as device code), fresh carry state, summary reset and
// the trajectory fingerprint of an existing full-mode output.
//
// None of this has a counterpart in the reference: its depth source is the SQS API
// (sqs/sqs.go:51) and its tests use one constant depth per test (main_test.go:288-304).
#include "ksa_internal.h"
#include "../../include/ksa_tracegen.h"

namespace {

constexpr int kBlock = 256;

__device__ __forceinline__ void gen_store(const ksa_gen_params& g, int A, long long v,
                                          unsigned long long h, long long* cell, long long ld) {
  if (A == 1) {
    cell[0] = v;
    return;
  }
  int64_t parts[3];
  ksa_gen_split3(&g, v, h, parts);
  if (A == 2) {
    cell[0] = parts[0];
    cell[ld] = (long long)((unsigned long long)parts[1] + (unsigned long long)parts[2]);
  } else {
    cell[0] = parts[0];
    cell[ld] = parts[1];
    cell[2 * ld] = parts[2];
  }
}

// Memoryless families: one thread per (tick-strip, trace); consecutive threads write consecutive
// traces of a row, so every store instruction of a warp covers 256 contiguous bytes.
__global__ void __launch_bounds__(kBlock) gen_iid_kernel(const __grid_constant__ ksa_gen_params g,
                                                         long long tick0, int T, int A,
                                                         long long trace0, long long N, long long ld,
                                                         long long* depth, int ticks_per_block) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  const unsigned long long key = ksa_trace_key(g.seed, 0, (unsigned long long)(trace0 + n));
  const int t_begin = blockIdx.y * ticks_per_block;
  const int t_end = min(T, t_begin + ticks_per_block);
  for (int t = t_begin; t < t_end; ++t) {
    const unsigned long long h = ksa_entry_word(key, (unsigned long long)(tick0 + 1 + t));
    const long long v = ksa_gen_iid_value(&g, h);
    gen_store(g, A, v, h, depth + (long long)t * A * ld + n, ld);
  }
}

// Random walk: sequential in time per trace, one thread per trace.
__global__ void __launch_bounds__(kBlock) gen_walk_kernel(const __grid_constant__ ksa_gen_params g,
                                                          long long tick0, int T, int A,
                                                          long long trace0, long long N, long long ld,
                                                          long long* depth, long long* walk) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  const unsigned long long key = ksa_trace_key(g.seed, 0, (unsigned long long)(trace0 + n));
  long long pos = walk[n];
  for (int t = 0; t < T; ++t) {
    const unsigned long long h = ksa_entry_word(key, (unsigned long long)(tick0 + 1 + t));
    pos = ksa_gen_walk_step(&g, pos, h);
    gen_store(g, A, pos, h, depth + (long long)t * A * ld + n, ld);
  }
  walk[n] = pos;
}

__global__ void __launch_bounds__(kBlock) walk_init_kernel(const __grid_constant__ ksa_gen_params g,
                                                           long long trace0, long long N,
                                                           long long* walk) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  walk[n] = ksa_gen_walk_start(&g, ksa_trace_key(g.seed, 0, (unsigned long long)(trace0 + n)));
}

__global__ void __launch_bounds__(kBlock) state_init_kernel(unsigned long long seed, long long trace0,
                                                            long long N, int lo, int hi, int* r,
                                                            long long* ju, long long* jd) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  r[n] = ksa_gen_replicas0(seed, (unsigned long long)(trace0 + n), lo, hi);
  ju[n] = 0;  // main.go:37 lastScaleUpTime := time.Now() at process start == tick 0
  jd[n] = 0;  // main.go:38
}

__global__ void __launch_bounds__(kBlock) summary_clear_kernel(ksa_summary* s, long long N) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  ksa_summary z;
  z.final_replicas = 0;
  z.min_replicas = INT_MAX;
  z.max_replicas = INT_MIN;
  z.up_applied = z.down_applied = z.up_noop = z.down_noop = z.up_cool = z.down_cool = 0u;
  z.poll_failed = z.up_failed = z.down_failed = 0u;
  z.sum_replicas = 0;
  z.hash = 0ull;
  s[n] = z;
}

// Synthetic failure mask (include/ksa_tracegen.h): one thread per trace and tick-strip, byte stores
// coalesced across traces.
__global__ void __launch_bounds__(kBlock) gen_faults_kernel(unsigned long long seed, unsigned poll_log2,
                                                            unsigned up_log2, unsigned down_log2,
                                                            long long tick0, int T, long long trace0,
                                                            long long N, long long ld, unsigned char* fail,
                                                            int ticks_per_block) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  const unsigned long long key = ksa_fault_key(seed, (unsigned long long)(trace0 + n));
  const int t_begin = blockIdx.y * ticks_per_block;
  const int t_end = min(T, t_begin + ticks_per_block);
  for (int t = t_begin; t < t_end; ++t)
    fail[(long long)t * ld + n] =
        ksa_gen_fault_byte(key, (unsigned long long)(tick0 + 1 + t), poll_log2, up_log2, down_log2);
}

// Fingerprint of an existing trajectory: one thread per trace, coalesced 4-byte / 1-byte reads,
// 8 ticks in flight per thread.
__global__ void __launch_bounds__(kBlock) traj_hash_kernel(const int* __restrict__ replicas,
                                                           const unsigned char* __restrict__ events,
                                                           int T, long long N, long long ld, int chain,
                                                           unsigned long long* hash) {
  const long long n = (long long)blockIdx.x * kBlock + threadIdx.x;
  if (n >= N) return;
  unsigned long long h = chain ? hash[n] : 0ull;
  constexpr int U = 8;
  int t = 0;
  for (; t + U <= T; t += U) {
    int r[U];
    unsigned e[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      r[j] = __ldg(replicas + (long long)(t + j) * ld + n);
      e[j] = __ldg(events + (long long)(t + j) * ld + n);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) h = ksa_hash_step(h, r[j], e[j]);
  }
  for (; t < T; ++t)
    h = ksa_hash_step(h, __ldg(replicas + (long long)t * ld + n), __ldg(events + (long long)t * ld + n));
  hash[n] = h;
}

inline unsigned blocks_for(long long n) { return (unsigned)((n + kBlock - 1) / kBlock); }

int check_gen(const ksa_gen_params* g, int64_t T, int32_t A, int64_t N, int64_t ld) {
  if (!g) return ksa_fail(KSA_EINVAL, "NULL ksa_gen_params");
  if (g->family < KSA_GEN_BANDS || g->family > KSA_GEN_ADVERSARIAL)
    return ksa_fail(KSA_EINVAL, "unknown generator family %d", g->family);
  if (g->reserved != 0 || g->reserved2 != 0) return ksa_fail(KSA_EINVAL, "reserved fields must be 0");
  if (A < 1 || A > KSA_MAX_ATTRS) return ksa_fail(KSA_EINVAL, "A=%d outside [1,3]", A);
  if (T < 0 || T > KSA_MAX_CHUNK_TICKS) return ksa_fail(KSA_EINVAL, "T=%lld outside [0, 2^30 - 64]", (long long)T);
  if (N < 0 || N > 0x7fffffffll || ld < N) return ksa_fail(KSA_EINVAL, "bad N=%lld / ld=%lld", (long long)N, (long long)ld);
  if (g->low_max < 0 || g->high_min <= g->low_max || g->high_max < g->high_min ||
      g->high_max > 0xFFFFFFFFll)
    return ksa_fail(KSA_EINVAL, "need 0 <= low_max < high_min <= high_max <= 2^32-1");
  if (g->family == KSA_GEN_WALK && (g->walk_step < 0 || g->walk_step > 0x7FFFFFFFll))
    return ksa_fail(KSA_EINVAL, "walk_step outside [0, 2^31-1]");
  if (g->overflow_log2 > 63) return ksa_fail(KSA_EINVAL, "overflow_log2 > 63");
  return KSA_OK;
}

}  // namespace

extern "C" int ksa_gen_traces(const ksa_gen_params* g, int64_t tick0, int64_t T, int32_t A,
                              int64_t trace0, int64_t N, int64_t ld, int64_t* depth_dev,
                              int64_t* walk_dev, void* stream) {
  int rc = check_gen(g, T, A, N, ld);
  if (rc != KSA_OK) return rc;
  if (!depth_dev) return ksa_fail(KSA_EINVAL, "NULL depth_dev");
  if (N == 0 || T == 0) return KSA_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (g->family == KSA_GEN_WALK) {
    if (!walk_dev) return ksa_fail(KSA_EINVAL, "KSA_GEN_WALK needs walk_dev");
    gen_walk_kernel<<<blocks_for(N), kBlock, 0, st>>>(*g, tick0, (int)T, A, trace0, N, ld,
                                                      (long long*)depth_dev, (long long*)walk_dev);
    KSA_LAUNCH_CHECK("gen_walk_kernel");
  } else {
    // Enough tick-strips to fill the GPU even for few traces; 65535 is the gridDim.y limit.
    int strips = (int)((T + 63) / 64);
    if (strips > 65535) strips = 65535;
    const int per = (int)((T + strips - 1) / strips);
    strips = (int)((T + per - 1) / per);
    gen_iid_kernel<<<dim3(blocks_for(N), (unsigned)strips), kBlock, 0, st>>>(
        *g, tick0, (int)T, A, trace0, N, ld, (long long*)depth_dev, per);
    KSA_LAUNCH_CHECK("gen_iid_kernel");
  }
  return KSA_OK;
}

extern "C" int ksa_gen_faults(uint64_t seed, uint32_t poll_log2, uint32_t up_log2, uint32_t down_log2,
                              int64_t tick0, int64_t T, int64_t trace0, int64_t N, int64_t ld,
                              uint8_t* fail_dev, void* stream) {
  if (!fail_dev) return ksa_fail(KSA_EINVAL, "NULL fail_dev");
  if (poll_log2 > 63 || up_log2 > 63 || down_log2 > 63) return ksa_fail(KSA_EINVAL, "log2 rates must be <= 63");
  if (T < 0 || T > KSA_MAX_CHUNK_TICKS || N < 0 || N > 0x7fffffffll || ld < N)
    return ksa_fail(KSA_EINVAL, "bad T/N/ld");
  if (N == 0 || T == 0) return KSA_OK;
  int strips = (int)((T + 63) / 64);
  if (strips > 65535) strips = 65535;
  const int per = (int)((T + strips - 1) / strips);
  strips = (int)((T + per - 1) / per);
  gen_faults_kernel<<<dim3(blocks_for(N), (unsigned)strips), kBlock, 0, (cudaStream_t)stream>>>(
      seed, poll_log2, up_log2, down_log2, tick0, (int)T, trace0, N, ld, fail_dev, per);
  KSA_LAUNCH_CHECK("gen_faults_kernel");
  return KSA_OK;
}

extern "C" int ksa_gen_walk_init(const ksa_gen_params* g, int64_t trace0, int64_t N,
                                 int64_t* walk_dev, void* stream) {
  int rc = check_gen(g, 0, 1, N, N);
  if (rc != KSA_OK) return rc;
  if (!walk_dev) return ksa_fail(KSA_EINVAL, "NULL walk_dev");
  if (N == 0) return KSA_OK;
  walk_init_kernel<<<blocks_for(N), kBlock, 0, (cudaStream_t)stream>>>(*g, trace0, N,
                                                                       (long long*)walk_dev);
  KSA_LAUNCH_CHECK("walk_init_kernel");
  return KSA_OK;
}

extern "C" int ksa_state_init(uint64_t seed, int64_t trace0, int64_t N, int32_t lo, int32_t hi,
                              const ksa_state* state, void* stream) {
  if (!state || !state->replicas_dev || !state->last_up_tick_dev || !state->last_down_tick_dev)
    return ksa_fail(KSA_EINVAL, "NULL state");
  if (N < 0 || hi < lo) return ksa_fail(KSA_EINVAL, "bad N or hi < lo");
  if (N == 0) return KSA_OK;
  state_init_kernel<<<blocks_for(N), kBlock, 0, (cudaStream_t)stream>>>(
      seed, trace0, N, lo, hi, state->replicas_dev, (long long*)state->last_up_tick_dev,
      (long long*)state->last_down_tick_dev);
  KSA_LAUNCH_CHECK("state_init_kernel");
  return KSA_OK;
}

extern "C" int ksa_summary_clear(ksa_summary* summary_dev, int64_t N, void* stream) {
  if (!summary_dev || N < 0) return ksa_fail(KSA_EINVAL, "NULL summary_dev or N < 0");
  if (N == 0) return KSA_OK;
  summary_clear_kernel<<<blocks_for(N), kBlock, 0, (cudaStream_t)stream>>>(summary_dev, N);
  KSA_LAUNCH_CHECK("summary_clear_kernel");
  return KSA_OK;
}

extern "C" int ksa_traj_hash(const int32_t* replicas_dev, const uint8_t* events_dev, int64_t T,
                             int64_t N, int64_t ld, int32_t chain, uint64_t* hash_dev, void* stream) {
  if (!replicas_dev || !events_dev || !hash_dev) return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (T < 0 || T > KSA_MAX_CHUNK_TICKS || N < 0 || ld < N) return ksa_fail(KSA_EINVAL, "bad T/N/ld");
  if (N == 0) return KSA_OK;
  traj_hash_kernel<<<blocks_for(N), kBlock, 0, (cudaStream_t)stream>>>(
      replicas_dev, events_dev, (int)T, N, ld, chain, (unsigned long long*)hash_dev);
  KSA_LAUNCH_CHECK("traj_hash_kernel");
  return KSA_OK;
}
