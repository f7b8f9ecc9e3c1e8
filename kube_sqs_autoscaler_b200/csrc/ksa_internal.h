// ksa_internal.h — shared by the translation units of libksareplay (not installed).
#pragma once

#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

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

// Time-parallel full replay for small trace counts (ksa_timepar.cu).
int ksa_timepar_plan(int64_t T, int64_t N, bool vector_ok, int* L_out, int* warm_out, int* wide_out);
int ksa_timepar_replay_full(const ksa_policy_k& pk, const int64_t* depth_dev, int64_t T, int32_t A, int64_t N,
                            int64_t ld, int64_t tick0, const ksa_state* state, int32_t* replicas_dev,
                            uint8_t* events_dev, int K, int L, int warm, int wide, cudaStream_t st);

static inline bool ksa_aligned(const void* p, size_t a) { return ((uintptr_t)p & (a - 1)) == 0; }
