// ksa_api.cu — the non-kernel half of the C ABI: error reporting, device helpers, the policy
// conversions the reference does implicitly in Go (int -> int32 truncation, Duration -> ticks), and
// the host-buffer pipeline that overlaps PCIe copies with the replay kernels.
#include "ksa_internal.h"

#include <cstring>
#include <new>

namespace {
thread_local char g_err[512] = "";
thread_local long long g_launches = 0;
thread_local int g_full_variant = -1;     // -1 = library default
thread_local int g_summary_variant = -1;
}  // namespace

bool ksa_replay_variant_exists(int mode, int id);  // ksa_replay.cu

int ksa_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
void ksa_count_launch(int n) { g_launches += n; }
int ksa_full_variant() { return g_full_variant; }
int ksa_summary_variant() { return g_summary_variant; }

// Go's int32(x) on a Go int: keep the low 32 bits (scale/scale.go:62,66-68,89,94-96).
static inline int go_int32(int64_t x) { return (int)(uint32_t)(uint64_t)x; }

// max(0, ceil(cool/poll)) without overflow.
static inline int64_t cool_ticks(int64_t cool, int64_t poll) {
  return cool <= 0 ? 0 : cool / poll + (cool % poll != 0 ? 1 : 0);
}

int ksa_policy_to_k(const ksa_policy* p, int32_t A, ksa_policy_k* out) {
  if (!p) return ksa_fail(KSA_EINVAL, "NULL policy");
  if (p->poll_period_ns <= 0)
    return ksa_fail(KSA_EINVAL, "poll_period_ns=%lld must be > 0 (main.go:41 sleeps one poll per tick)",
                    (long long)p->poll_period_ns);
  if (p->reserved != 0) return ksa_fail(KSA_EINVAL, "ksa_policy.reserved must be 0");
  if (A < 1 || A > KSA_MAX_ATTRS) return ksa_fail(KSA_EINVAL, "A=%d outside [1,3]", A);
  if ((p->attr_mask & ~7u) != 0 || __builtin_popcount(p->attr_mask) != A)
    return ksa_fail(KSA_EINVAL, "attr_mask=0x%x selects %d attributes but A=%d planes are stored",
                    p->attr_mask, __builtin_popcount(p->attr_mask & 7u), A);
  out->up_msgs = p->scale_up_messages;
  out->down_msgs = p->scale_down_messages;
  out->cu = cool_ticks(p->scale_up_cool_down_ns, p->poll_period_ns);
  out->cd = cool_ticks(p->scale_down_cool_down_ns, p->poll_period_ns);
  out->max32 = go_int32(p->max_pods);
  out->min32 = go_int32(p->min_pods);
  out->up32 = go_int32(p->scale_up_pods);
  out->down32 = go_int32(p->scale_down_pods);
  return KSA_OK;
}

extern "C" int ksa_abi_version(void) { return KSA_ABI_VERSION; }
extern "C" const char* ksa_last_error(void) { return g_err; }

extern "C" int ksa_set_device(int device) {
  KSA_CUDA(cudaSetDevice(device));
  return KSA_OK;
}
extern "C" int ksa_device_count(void) {
  int n = 0;
  KSA_CUDA(cudaGetDeviceCount(&n));
  return n;
}
extern "C" int ksa_dev_alloc(void** ptr_dev, size_t bytes) {
  if (!ptr_dev) return ksa_fail(KSA_EINVAL, "NULL out pointer");
  cudaError_t e = cudaMalloc(ptr_dev, bytes);
  if (e == cudaErrorMemoryAllocation) {
    cudaGetLastError();
    return ksa_fail(KSA_ENOMEM, "cudaMalloc(%zu) out of device memory", bytes);
  }
  KSA_CUDA(e);
  return KSA_OK;
}
extern "C" int ksa_dev_free(void* ptr_dev) {
  KSA_CUDA(cudaFree(ptr_dev));
  return KSA_OK;
}
extern "C" int ksa_host_alloc(void** ptr_host, size_t bytes) {
  if (!ptr_host) return ksa_fail(KSA_EINVAL, "NULL out pointer");
  cudaError_t e = cudaMallocHost(ptr_host, bytes);
  if (e == cudaErrorMemoryAllocation) {
    cudaGetLastError();
    return ksa_fail(KSA_ENOMEM, "cudaMallocHost(%zu) failed", bytes);
  }
  KSA_CUDA(e);
  return KSA_OK;
}
extern "C" int ksa_host_free(void* ptr_host) {
  KSA_CUDA(cudaFreeHost(ptr_host));
  return KSA_OK;
}
extern "C" int ksa_host_register(void* ptr_host, size_t bytes) {
  if (!ptr_host || bytes == 0) return ksa_fail(KSA_EINVAL, "ksa_host_register: NULL pointer or no bytes");
  cudaError_t e = cudaHostRegister(ptr_host, bytes, cudaHostRegisterPortable);
  if (e == cudaErrorMemoryAllocation) {
    cudaGetLastError();
    return ksa_fail(KSA_ENOMEM, "cudaHostRegister(%zu) failed", bytes);
  }
  KSA_CUDA(e);
  return KSA_OK;
}
extern "C" int ksa_host_unregister(void* ptr_host) {
  KSA_CUDA(cudaHostUnregister(ptr_host));
  return KSA_OK;
}
extern "C" int ksa_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream) {
  KSA_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return KSA_OK;
}
extern "C" int ksa_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream) {
  KSA_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return KSA_OK;
}
extern "C" int ksa_stream_sync(void* stream) {
  KSA_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return KSA_OK;
}

extern "C" int ksa_policy_default(ksa_policy* p) {
  if (!p) return ksa_fail(KSA_EINVAL, "NULL policy");
  p->scale_up_messages = 100;                    // main.go:86
  p->scale_down_messages = 10;                   // main.go:87
  p->max_pods = 5;                               // main.go:90
  p->min_pods = 1;                               // main.go:91
  p->scale_up_pods = 1;                          // main.go:88
  p->scale_down_pods = 1;                        // main.go:89
  p->poll_period_ns = 5ll * 1000000000ll;        // main.go:83
  p->scale_up_cool_down_ns = 10ll * 1000000000ll;   // main.go:85
  p->scale_down_cool_down_ns = 30ll * 1000000000ll; // main.go:84
  p->attr_mask = KSA_ATTR_MESSAGES | KSA_ATTR_MESSAGES_DELAYED | KSA_ATTR_MESSAGES_NOT_VISIBLE;  // main.go:28,93
  p->reserved = 0;
  return KSA_OK;
}

extern "C" int ksa_policy_cool_ticks(const ksa_policy* p, int64_t* up_ticks, int64_t* down_ticks) {
  if (!p || !up_ticks || !down_ticks) return ksa_fail(KSA_EINVAL, "NULL pointer");
  if (p->poll_period_ns <= 0) return ksa_fail(KSA_EINVAL, "poll_period_ns must be > 0");
  *up_ticks = cool_ticks(p->scale_up_cool_down_ns, p->poll_period_ns);
  *down_ticks = cool_ticks(p->scale_down_cool_down_ns, p->poll_period_ns);
  return KSA_OK;
}

extern "C" int ksa_set_variant(int32_t full_variant, int32_t summary_variant) {
  if (full_variant > 0 && !ksa_replay_variant_exists(0, full_variant))
    return ksa_fail(KSA_EINVAL, "unknown full-mode variant %d", full_variant);
  if (summary_variant > 0 && !ksa_replay_variant_exists(1, summary_variant))
    return ksa_fail(KSA_EINVAL, "unknown summary-mode variant %d", summary_variant);
  // 0 = library default; the scalar kernels (table id 0) are selected with -2.
  g_full_variant = full_variant == 0 ? -1 : (full_variant == -2 ? 0 : full_variant);
  g_summary_variant = summary_variant == 0 ? -1 : (summary_variant == -2 ? 0 : summary_variant);
  return KSA_OK;
}

extern "C" int64_t ksa_launch_count(void) {
  const long long n = g_launches;
  g_launches = 0;
  return n;
}

// ---- host-buffer pipeline -----------------------------------------------------------------------

namespace {
// The pipeline calls work on the pipeline's device and hand the caller's current device back.
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != device) cudaSetDevice(device);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};
}  // namespace

struct ksa_pipeline {
  int device;
  int64_t N, ld, chunk;
  int32_t A;
  cudaStream_t s_in, s_k, s_out;
  cudaEvent_t in_done[2], k_done[2], out_done[2];
  long long* depth[2];
  int* replicas[2];
  unsigned char* events[2];
  unsigned char* fail[2];  // failure-mask staging (ksa_pipeline_replay_faults)
  int* st_r[2];            // [0] full-mode state, [1] summary-mode state (same start values)
  long long* st_ju[2];
  long long* st_jd[2];
  ksa_summary* summary;
};

static void pipeline_free(ksa_pipeline* p) {
  if (!p) return;
  for (int i = 0; i < 2; ++i) {
    cudaFree(p->depth[i]);
    cudaFree(p->replicas[i]);
    cudaFree(p->events[i]);
    cudaFree(p->fail[i]);
    cudaFree(p->st_r[i]);
    cudaFree(p->st_ju[i]);
    cudaFree(p->st_jd[i]);
    if (p->in_done[i]) cudaEventDestroy(p->in_done[i]);
    if (p->k_done[i]) cudaEventDestroy(p->k_done[i]);
    if (p->out_done[i]) cudaEventDestroy(p->out_done[i]);
  }
  cudaFree(p->summary);
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_k) cudaStreamDestroy(p->s_k);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  delete p;
}

#define KSA_PIPE_CUDA(call)                                                                  \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      pipeline_free(p);                                                                      \
      cudaGetLastError();                                                                    \
      return ksa_fail(e_ == cudaErrorMemoryAllocation ? KSA_ENOMEM : KSA_ECUDA, "%s: %s", #call, \
                      cudaGetErrorString(e_));                                               \
    }                                                                                        \
  } while (0)

extern "C" int ksa_pipeline_create(ksa_pipeline** out, int device, int64_t N, int32_t A,
                                   int64_t chunk_ticks) {
  if (!out) return ksa_fail(KSA_EINVAL, "NULL out pointer");
  *out = nullptr;
  if (N < 1 || N > 0x7fffffffll) return ksa_fail(KSA_EINVAL, "N=%lld outside [1, 2^31-1]", (long long)N);
  if (A < 1 || A > KSA_MAX_ATTRS) return ksa_fail(KSA_EINVAL, "A=%d outside [1,3]", A);
  if (chunk_ticks < 1 || chunk_ticks > KSA_MAX_CHUNK_TICKS)
    return ksa_fail(KSA_EINVAL, "chunk_ticks=%lld outside [1, 2^30 - 64]", (long long)chunk_ticks);
  ksa_pipeline* p = new (std::nothrow) ksa_pipeline();
  if (!p) return ksa_fail(KSA_ENOMEM, "out of host memory");
  std::memset(p, 0, sizeof(*p));
  p->device = device;
  p->N = N;
  p->ld = (N + 15) / 16 * 16;  // padded pitch so the vector kernels apply to any N
  p->chunk = chunk_ticks;
  p->A = A;
  DeviceGuard guard(device);
  KSA_PIPE_CUDA(cudaGetLastError());
  KSA_PIPE_CUDA(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
  KSA_PIPE_CUDA(cudaStreamCreateWithFlags(&p->s_k, cudaStreamNonBlocking));
  KSA_PIPE_CUDA(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
  const size_t cells = (size_t)chunk_ticks * (size_t)p->ld;
  for (int i = 0; i < 2; ++i) {
    KSA_PIPE_CUDA(cudaEventCreateWithFlags(&p->in_done[i], cudaEventDisableTiming));
    KSA_PIPE_CUDA(cudaEventCreateWithFlags(&p->k_done[i], cudaEventDisableTiming));
    KSA_PIPE_CUDA(cudaEventCreateWithFlags(&p->out_done[i], cudaEventDisableTiming));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->depth[i], cells * (size_t)A * sizeof(long long)));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->replicas[i], cells * sizeof(int)));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->events[i], cells));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->fail[i], cells));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->st_r[i], (size_t)p->ld * sizeof(int)));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->st_ju[i], (size_t)p->ld * sizeof(long long)));
    KSA_PIPE_CUDA(cudaMalloc((void**)&p->st_jd[i], (size_t)p->ld * sizeof(long long)));
  }
  KSA_PIPE_CUDA(cudaMalloc((void**)&p->summary, (size_t)p->ld * sizeof(ksa_summary)));
  *out = p;
  return KSA_OK;
}

extern "C" int ksa_pipeline_destroy(ksa_pipeline* p) {
  if (!p) return KSA_OK;
  DeviceGuard guard(p->device);
  cudaDeviceSynchronize();
  pipeline_free(p);
  return KSA_OK;
}

// Row-pitch-converting async copies: host rows are tight (N elements), device rows are padded (ld).
static cudaError_t copy_rows_h2d(void* dst, const void* src, size_t elem, int64_t N, int64_t ld,
                                 int64_t rows, cudaStream_t st) {
  if (N == ld) return cudaMemcpyAsync(dst, src, elem * (size_t)N * (size_t)rows, cudaMemcpyHostToDevice, st);
  return cudaMemcpy2DAsync(dst, elem * (size_t)ld, src, elem * (size_t)N, elem * (size_t)N,
                           (size_t)rows, cudaMemcpyHostToDevice, st);
}
static cudaError_t copy_rows_d2h(void* dst, const void* src, size_t elem, int64_t N, int64_t ld,
                                 int64_t rows, cudaStream_t st) {
  if (N == ld) return cudaMemcpyAsync(dst, src, elem * (size_t)N * (size_t)rows, cudaMemcpyDeviceToHost, st);
  return cudaMemcpy2DAsync(dst, elem * (size_t)N, src, elem * (size_t)ld, elem * (size_t)N,
                           (size_t)rows, cudaMemcpyDeviceToHost, st);
}

// fail_host == NULL: the plain replay; otherwise the *_faults kernels with the mask staged
// alongside the depths (one more byte per decision on the H2D stream).
static int pipeline_run_inner(ksa_pipeline* p, const ksa_policy* policy, const int64_t* depth_host,
                        const uint8_t* fail_host, int64_t T, int64_t N, int64_t tick0,
                        int32_t* state_replicas_host, int64_t* state_last_up_host,
                        int64_t* state_last_down_host, int32_t* replicas_host, uint8_t* events_host,
                        ksa_summary* summary_host) {
  if (!p || !policy || !depth_host || !state_replicas_host || !state_last_up_host || !state_last_down_host)
    return ksa_fail(KSA_EINVAL, "NULL pipeline, policy, depth or state");
  if (N < 1 || N > p->N) return ksa_fail(KSA_ELIMIT, "N=%lld exceeds the pipeline's %lld traces", (long long)N, (long long)p->N);
  if (T < 1) return ksa_fail(KSA_EINVAL, "T=%lld < 1", (long long)T);
  if (tick0 < 0) return ksa_fail(KSA_EINVAL, "tick0 < 0");
  ksa_policy_k probe;
  int rc = ksa_policy_to_k(policy, p->A, &probe);  // validate before touching the device
  if (rc != KSA_OK) return rc;
  DeviceGuard guard(p->device);
  const int64_t ld = p->ld;
  const int32_t A = p->A;
  const bool want_full = replicas_host || events_host;
  // With no per-tick output requested the summary pass runs anyway: it advances the carry state.
  const bool want_summary = summary_host != nullptr || !want_full;
  const bool advance_with_summary = !want_full;

  // Carry state in (both copies when both passes run), summary reset.
  for (int i = 0; i < 2; ++i) {
    if (i == 0 && !want_full) continue;
    if (i == 1 && !want_summary) continue;
    KSA_CUDA(cudaMemcpyAsync(p->st_r[i], state_replicas_host, (size_t)N * sizeof(int), cudaMemcpyHostToDevice, p->s_k));
    KSA_CUDA(cudaMemcpyAsync(p->st_ju[i], state_last_up_host, (size_t)N * sizeof(long long), cudaMemcpyHostToDevice, p->s_k));
    KSA_CUDA(cudaMemcpyAsync(p->st_jd[i], state_last_down_host, (size_t)N * sizeof(long long), cudaMemcpyHostToDevice, p->s_k));
  }
  if (want_summary) {
    rc = ksa_summary_clear(p->summary, N, p->s_k);
    if (rc != KSA_OK) return rc;
  }
  const int state_full = 0, state_sum = 1;
  ksa_state sf = {p->st_r[state_full], (int64_t*)p->st_ju[state_full], (int64_t*)p->st_jd[state_full]};
  ksa_state ss = {p->st_r[state_sum], (int64_t*)p->st_ju[state_sum], (int64_t*)p->st_jd[state_sum]};

  const int64_t chunks = (T + p->chunk - 1) / p->chunk;
  for (int64_t c = 0; c < chunks; ++c) {
    const int s = (int)(c & 1);
    const int64_t t_begin = c * p->chunk;
    const int64_t tc = (T - t_begin) < p->chunk ? (T - t_begin) : p->chunk;
    // stage s's depth buffer is free once the kernels of chunk c-2 have consumed it
    if (c >= 2) KSA_CUDA(cudaStreamWaitEvent(p->s_in, p->k_done[s], 0));
    KSA_CUDA(copy_rows_h2d(p->depth[s], depth_host + (size_t)t_begin * (size_t)A * (size_t)N,
                           sizeof(long long), N, ld, tc * A, p->s_in));
    if (fail_host)
      KSA_CUDA(copy_rows_h2d(p->fail[s], fail_host + (size_t)t_begin * (size_t)N, 1, N, ld, tc, p->s_in));
    KSA_CUDA(cudaEventRecord(p->in_done[s], p->s_in));
    KSA_CUDA(cudaStreamWaitEvent(p->s_k, p->in_done[s], 0));
    // stage s's output buffers are free once chunk c-2's D2H copies are done
    if (c >= 2 && want_full) KSA_CUDA(cudaStreamWaitEvent(p->s_k, p->out_done[s], 0));
    if (want_full) {
      int32_t* r_dev = replicas_host ? p->replicas[s] : nullptr;
      uint8_t* e_dev = events_host ? p->events[s] : nullptr;
      rc = fail_host ? ksa_replay_full_faults(policy, (const int64_t*)p->depth[s], p->fail[s], tc, A, N, ld,
                                              tick0 + t_begin, &sf, r_dev, e_dev, p->s_k)
                     : ksa_replay_full(policy, (const int64_t*)p->depth[s], tc, A, N, ld, tick0 + t_begin, &sf,
                                       r_dev, e_dev, p->s_k);
      if (rc != KSA_OK) return rc;
    }
    if (want_summary) {
      rc = fail_host ? ksa_replay_summary_faults(policy, (const int64_t*)p->depth[s], p->fail[s], tc, A, N, ld,
                                                 tick0 + t_begin, &ss, p->summary, p->s_k)
                     : ksa_replay_summary(policy, (const int64_t*)p->depth[s], tc, A, N, ld, tick0 + t_begin,
                                          &ss, p->summary, p->s_k);
      if (rc != KSA_OK) return rc;
    }
    KSA_CUDA(cudaEventRecord(p->k_done[s], p->s_k));
    if (want_full) {
      KSA_CUDA(cudaStreamWaitEvent(p->s_out, p->k_done[s], 0));
      if (replicas_host)
        KSA_CUDA(copy_rows_d2h(replicas_host + (size_t)t_begin * (size_t)N, p->replicas[s], sizeof(int), N, ld, tc, p->s_out));
      if (events_host)
        KSA_CUDA(copy_rows_d2h(events_host + (size_t)t_begin * (size_t)N, p->events[s], 1, N, ld, tc, p->s_out));
      KSA_CUDA(cudaEventRecord(p->out_done[s], p->s_out));
    }
  }
  // Carry state and summaries out (ordered after the last kernel on s_k).
  const int adv = advance_with_summary ? state_sum : state_full;
  KSA_CUDA(cudaMemcpyAsync(state_replicas_host, p->st_r[adv], (size_t)N * sizeof(int), cudaMemcpyDeviceToHost, p->s_k));
  KSA_CUDA(cudaMemcpyAsync(state_last_up_host, p->st_ju[adv], (size_t)N * sizeof(long long), cudaMemcpyDeviceToHost, p->s_k));
  KSA_CUDA(cudaMemcpyAsync(state_last_down_host, p->st_jd[adv], (size_t)N * sizeof(long long), cudaMemcpyDeviceToHost, p->s_k));
  if (summary_host)
    KSA_CUDA(cudaMemcpyAsync(summary_host, p->summary, (size_t)N * sizeof(ksa_summary), cudaMemcpyDeviceToHost, p->s_k));
  KSA_CUDA(cudaStreamSynchronize(p->s_in));
  KSA_CUDA(cudaStreamSynchronize(p->s_k));
  KSA_CUDA(cudaStreamSynchronize(p->s_out));
  return KSA_OK;
}

// The calls are synchronous: when they return, nothing may still be reading the caller's host buffers
// or writing into them, also after an error. pipeline_run_inner returns at the first failing call;
// this wrapper then waits for whatever had been queued on the three streams (ignoring further
// errors) before the status goes back, so the caller can free or reuse its buffers and the next
// call finds the staging buffers and events idle.
static int pipeline_run(ksa_pipeline* p, const ksa_policy* policy, const int64_t* depth_host,
                        const uint8_t* fail_host, int64_t T, int64_t N, int64_t tick0,
                        int32_t* state_replicas_host, int64_t* state_last_up_host,
                        int64_t* state_last_down_host, int32_t* replicas_host, uint8_t* events_host,
                        ksa_summary* summary_host) {
  const int rc = pipeline_run_inner(p, policy, depth_host, fail_host, T, N, tick0, state_replicas_host,
                                    state_last_up_host, state_last_down_host, replicas_host, events_host,
                                    summary_host);
  if (rc != KSA_OK && p) {
    DeviceGuard guard(p->device);
    cudaStreamSynchronize(p->s_in);
    cudaStreamSynchronize(p->s_k);
    cudaStreamSynchronize(p->s_out);
    cudaGetLastError();  // the message of the first failure is already in ksa_last_error()
  }
  return rc;
}

extern "C" int ksa_pipeline_replay(ksa_pipeline* p, const ksa_policy* policy,
                                   const int64_t* depth_host, int64_t T, int64_t N, int64_t tick0,
                                   int32_t* state_replicas_host, int64_t* state_last_up_host,
                                   int64_t* state_last_down_host, int32_t* replicas_host,
                                   uint8_t* events_host, ksa_summary* summary_host) {
  return pipeline_run(p, policy, depth_host, nullptr, T, N, tick0, state_replicas_host, state_last_up_host,
                      state_last_down_host, replicas_host, events_host, summary_host);
}

extern "C" int ksa_pipeline_replay_faults(ksa_pipeline* p, const ksa_policy* policy,
                                          const int64_t* depth_host, const uint8_t* fail_host, int64_t T,
                                          int64_t N, int64_t tick0, int32_t* state_replicas_host,
                                          int64_t* state_last_up_host, int64_t* state_last_down_host,
                                          int32_t* replicas_host, uint8_t* events_host,
                                          ksa_summary* summary_host) {
  if (!fail_host) return ksa_fail(KSA_EINVAL, "NULL fail_host (use ksa_pipeline_replay)");
  return pipeline_run(p, policy, depth_host, fail_host, T, N, tick0, state_replicas_host, state_last_up_host,
                      state_last_down_host, replicas_host, events_host, summary_host);
}
