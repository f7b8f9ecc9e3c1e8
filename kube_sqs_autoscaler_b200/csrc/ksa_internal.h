// ksa_internal.h — shared by the translation units of libksareplay (not installed).
#pragma once

#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <functional>

#include "../../include/ksa_replay.h"
#include "ksa_policy.cuh"

// Error plumbing (ksa_api.cu): formats the thread-local message and returns `code`.
int ksa_fail(int code, const char* fmt, ...);
// Launch accounting for ksa_launch_count().
void ksa_count_launch(int n = 1);

#define KSA_CUDA(call)                                                                    \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return ksa_fail(KSA_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                      __FILE__, __LINE__);                                                \
  } while (0)

// After a kernel launch: surface launch-configuration errors without synchronising.
#define KSA_LAUNCH_CHECK(name)                                                             \
  do {                                                                                     \
    cudaError_t e_ = cudaGetLastError();                                                   \
    if (e_ != cudaSuccess)                                                                 \
      return ksa_fail(KSA_ECUDA, "launch of %s failed: %s", name, cudaGetErrorString(e_)); \
    ksa_count_launch();                                                                    \
  } while (0)

// Host: validate a ksa_policy and convert it to the kernel-side record. `A` is the number of
// attribute planes the caller stores; popcount(attr_mask) must match it.
int ksa_policy_to_k(const ksa_policy* p, int32_t A, ksa_policy_k* out);

// Variant selection (ksa_set_variant).
int ksa_full_variant();
int ksa_summary_variant();

// Time-parallel replays for small trace counts (ksa_timepar.cu); fail_dev == NULL: no failure mask.
// `plain(unresolved_dev)` launches the one-trace-per-thread kernel of the call for the traces whose
// flag in unresolved_dev[N] is set (the ones the chunked passes gave up); it runs last, on the same stream.
typedef std::function<int(const int* unresolved_dev)> ksa_plain_launch;
int ksa_timepar_plan(int64_t T, int64_t N, bool summary, bool faults, bool vector_ok, int* L_out, int* warm_out,
                     int* wide_out);
int ksa_timepar_replay_full(const ksa_policy_k& pk, const int64_t* depth_dev, const uint8_t* fail_dev, int64_t T,
                            int32_t A, int64_t N, int64_t ld, int64_t tick0, const ksa_state* state,
                            int32_t* replicas_dev, uint8_t* events_dev, int K, int L, int warm, int wide,
                            const ksa_plain_launch& plain, cudaStream_t st);
int ksa_timepar_replay_summary(const ksa_policy_k& pk, bool ordered, const int64_t* depth_dev, const uint8_t* fail_dev,
                               int64_t T, int32_t A, int64_t N, int64_t ld, int64_t tick0, const ksa_state* state,
                               ksa_summary* summary_dev, int K, int L, int warm, const ksa_plain_launch& plain,
                               cudaStream_t st);
// Its pass 1: the default summary kernel over (trace blocks, chunks of L ticks) (ksa_replay.cu).
int ksa_replay_summary_chunks(const ksa_policy_k& pk, bool ordered, const int64_t* depth_dev, const uint8_t* fail_dev,
                              int L, int64_t T, int32_t A, int64_t N, int64_t ld, int64_t tick0, int K, int32_t* st_r,
                              int64_t* st_ju, int64_t* st_jd, ksa_summary* partial, const int* marks_dev, int mark,
                              cudaStream_t stream);

static inline bool ksa_aligned(const void* p, size_t a) { return ((uintptr_t)p & (a - 1)) == 0; }
