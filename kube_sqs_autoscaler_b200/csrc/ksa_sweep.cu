// ksa_sweep.cu — ksa_replay_sweep: the same N traces replayed under P policies (the parameter grid
// of main.go:83-91), one ksa_summary per (policy, trace).
//
// Thread = one trace x a register tile of R policies: each depth value is loaded once and fed to
// R independent state machines (ksa_policy.cuh), so the depth block is re-read P/R times in total
// and stays resident in the 126 MB L2. Unlike the single-policy replay this kernel is bound by
// integer issue, not HBM: bytes per decision are ~0.01.
#include "ksa_internal.h"

#include <vector>

namespace {

struct SweepArgs {
  const ksa_policy_k* policies;  // [P] device
  const long long* depth;
  long long ld;
  long long N;
  int P, T, A;
  const int* replicas0;
  ksa_summary* summary;  // [P][N]
};

template <int R, int D, int BLOCK>
__global__ void __launch_bounds__(BLOCK) ksa_sweep_kernel(const __grid_constant__ SweepArgs a) {
  const long long n = (long long)blockIdx.x * BLOCK + threadIdx.x;
  if (n >= a.N) return;
  const int p0 = blockIdx.y * R;
  const int T = a.T, A = a.A;

  // The R policy records of this block: block-uniform loads (served from L1/constant path).
  ksa_policy_k pk[R];
#pragma unroll
  for (int q = 0; q < R; ++q) pk[q] = a.policies[min(p0 + q, a.P - 1)];

  ksa_trace_regs s[R];
  int rmin[R], rmax[R];
  unsigned cnt[R][6];
  long long sum[R];
  unsigned long long hash[R];
  const int r0 = a.replicas0[n];
#pragma unroll
  for (int q = 0; q < R; ++q) {
    const ksa_step_consts c(pk[q]);
    s[q].r = r0;
    s[q].lu = ksa_local_stamp(pk[q].cu, c.cu, 0, 0);  // fresh process: both stamps at tick 0,
    s[q].ld = ksa_local_stamp(pk[q].cd, c.cd, 0, 0);  // main.go:37-38
    rmin[q] = INT_MAX;
    rmax[q] = INT_MIN;
#pragma unroll
    for (int b = 0; b < 6; ++b) cnt[q][b] = 0u;
    sum[q] = 0;
    hash[q] = 0ull;
  }

  const long long row = (long long)A * a.ld;
  const long long* dp = a.depth + n;
  auto load_sum = [&](int t) {
    unsigned long long acc = (unsigned long long)__ldg(dp + (long long)t * row);
    for (int p = 1; p < A; ++p) acc += (unsigned long long)__ldg(dp + (long long)t * row + p * a.ld);
    return (long long)acc;  // sqs.go:56-64, wrapping
  };
  long long buf[D];
#pragma unroll
  for (int j = 0; j < D; ++j) buf[j] = j < T ? load_sum(j) : 0;

#pragma unroll 1
  for (int t = 0; t < T; t += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const int tt = t + j;
      if (tt < T) {
        const long long msgs = buf[j];
        if (tt + D < T) buf[j] = load_sum(tt + D);
        const int i = tt + 1;
#pragma unroll
        for (int q = 0; q < R; ++q) {
          const ksa_step_consts c(pk[q]);
          const unsigned ev = ksa_step(c, msgs, i, i - c.cu, i - c.cd, s[q]);
          const int r = s[q].r;
          rmin[q] = min(rmin[q], r);
          rmax[q] = max(rmax[q], r);
#pragma unroll
          for (int b = 0; b < 6; ++b) cnt[q][b] += (ev >> b) & 1u;
          sum[q] = (long long)((unsigned long long)sum[q] + (unsigned long long)(long long)r);
          hash[q] = ksa_hash_step(hash[q], r, ev);
        }
      }
    }
  }

#pragma unroll
  for (int q = 0; q < R; ++q) {
    if (p0 + q < a.P) {
      ksa_summary z;
      z.final_replicas = s[q].r;
      z.min_replicas = rmin[q];
      z.max_replicas = rmax[q];
      z.up_applied = cnt[q][0];
      z.down_applied = cnt[q][1];
      z.up_noop = cnt[q][2];
      z.down_noop = cnt[q][3];
      z.up_cool = cnt[q][4];
      z.down_cool = cnt[q][5];
      z.reserved = 0u;
      z.ticks = T;
      z.sum_replicas = sum[q];
      z.hash = hash[q];
      a.summary[(long long)(p0 + q) * a.N + n] = z;
    }
  }
}

}  // namespace

extern "C" int ksa_replay_sweep(const ksa_policy* policies, int32_t P, void* policies_dev,
                                const int64_t* depth_dev, int64_t T, int32_t A, int64_t N, int64_t ld,
                                const int32_t* replicas0_dev, ksa_summary* summary_dev, void* stream) {
  if (!policies || !policies_dev || !depth_dev || !replicas0_dev || !summary_dev)
    return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (P < 1 || P > (1 << 20)) return ksa_fail(KSA_EINVAL, "P=%d outside [1, 2^20]", P);
  if (T < 1 || T > KSA_MAX_CHUNK_TICKS) return ksa_fail(KSA_EINVAL, "T=%lld outside [1, 2^30]", (long long)T);
  if (N < 1 || N > 0x7fffffffll || ld < N) return ksa_fail(KSA_EINVAL, "bad N=%lld / ld=%lld", (long long)N, (long long)ld);
  // Convert on the host, then upload. The source is pageable memory, for which cudaMemcpyAsync
  // returns only after the bytes have been staged, so the vector may die at function exit.
  std::vector<ksa_policy_k> host_table((size_t)P);
  for (int q = 0; q < P; ++q) {
    int rc = ksa_policy_to_k(&policies[q], A, &host_table[q]);
    if (rc != KSA_OK) return rc;
    if (policies[q].attr_mask != policies[0].attr_mask)
      return ksa_fail(KSA_EINVAL, "policy %d: attr_mask differs from policy 0", q);
  }
  cudaStream_t st = (cudaStream_t)stream;
  KSA_CUDA(cudaMemcpyAsync(policies_dev, host_table.data(), sizeof(ksa_policy_k) * (size_t)P,
                           cudaMemcpyHostToDevice, st));
  SweepArgs a;
  a.policies = (const ksa_policy_k*)policies_dev;
  a.depth = (const long long*)depth_dev;
  a.ld = ld;
  a.N = N;
  a.P = P;
  a.T = (int)T;
  a.A = A;
  a.replicas0 = replicas0_dev;
  a.summary = summary_dev;
  constexpr int R = 4, D = 4, BLOCK = 128;
  dim3 grid((unsigned)((N + BLOCK - 1) / BLOCK), (unsigned)((P + R - 1) / R));
  if (grid.y > 65535) return ksa_fail(KSA_ELIMIT, "P too large for one launch");
  ksa_sweep_kernel<R, D, BLOCK><<<grid, BLOCK, 0, st>>>(a);
  KSA_LAUNCH_CHECK("ksa_sweep_kernel");
  return KSA_OK;
}
