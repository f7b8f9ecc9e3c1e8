/*
 * ksa_replay.h — C ABI of libksareplay, the B200 (sm_100a) batched replay engine for the
 * kube-sqs-autoscaler scaling policy.
 *
 * The reference (Go, /root/reference) has no FFI; its interface for this path is in-process:
 *   Run(p *scale.PodAutoScaler, sqs *kubesqs.SqsClient)        main.go:35
 *   (*SqsClient).NumMessages() (int, error)                    sqs/sqs.go:45
 *   (*PodAutoScaler).ScaleUp(ctx) error / ScaleDown(ctx) error scale/scale.go:54,81
 * parameterised by the flag globals main.go:16-33,83-91 and the PodAutoScaler struct
 * scale/scale.go:21-29. The entry points below are what a cgo binding for a batched replay of
 * that path binds (INTEGRATION.md shows the Go side): the SQS responses become an int64
 * depth array [T][A][N], the Deployment's Spec.Replicas becomes an int32 register per trace.
 *
 * Conventions
 *  - Plain C types only. All `*_dev` pointers are CUDA device pointers on the CURRENT device of
 *    the calling thread (cudaSetDevice / ksa_set_device); `stream` is a cudaStream_t passed as
 *    void* (NULL = legacy default stream). All launches are asynchronous on that stream.
 *  - Every function returns 0 (KSA_OK) or a negative KSA_E* code; ksa_last_error() returns the
 *    thread-local message of the last failure. Nothing aborts, nothing falls back to the CPU.
 *  - The caller owns every buffer. The replay entry points allocate nothing, on the device or on
 *    the host heap (ksa_dev_alloc / ksa_host_alloc and ksa_pipeline_create are the allocating
 *    helpers). Two fixed arrays live in the library's device image on every GPU that loads it:
 *    the policy table of ksa_replay_sweep_static (192 KB) and the scratch records of the
 *    time-parallel paths of the four replay entry points (76 MB, see ksa_replay_full).
 *  - Ticks: tick k (1-based, absolute) happens at t0 + k*poll (main.go:41). A call replays the
 *    T ticks tick0+1 .. tick0+T; `tick0` is the number of ticks already replayed for these
 *    traces (0 for a fresh start), which is what lets long traces be streamed in T-chunks.
 *  - Row layout: element (t, a, n) of depth is depth[(t*A + a)*ld + n]; replicas and events rows
 *    use the same pitch: replicas[t*ld + n], events[t*ld + n]. ld >= N, in ELEMENTS.
 *    Fast (vectorised) path: ld % 16 == 0 and all base pointers 32-byte aligned (cudaMalloc gives
 *    256); anything else silently takes the scalar kernel (same results, lower bandwidth).
 */
#ifndef KSA_REPLAY_H
#define KSA_REPLAY_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KSA_ABI_VERSION 1

/* ---- status codes --------------------------------------------------------------------- */
#define KSA_OK 0
#define KSA_EINVAL (-1)  /* bad argument (poll<=0, T/N/ld/A out of range, NULL pointer, ...)   */
#define KSA_ECUDA (-2)   /* a CUDA runtime call failed; message carries cudaGetErrorString     */
#define KSA_ENOMEM (-3)  /* an allocating helper (ksa_dev_alloc, pipeline create) failed        */
#define KSA_ELIMIT (-4)  /* a documented size limit was exceeded                                */

/* ---- event bits: one per log line of the reference ------------------------------------- */
#define KSA_EV_UP_APPLIED 0x01   /* "Scale up successful"               scale/scale.go:77  */
#define KSA_EV_DOWN_APPLIED 0x02 /* "Scale down successful"             scale/scale.go:105 */
#define KSA_EV_UP_NOOP 0x04      /* "More than max pods running"        scale/scale.go:63  */
#define KSA_EV_DOWN_NOOP 0x08    /* "Less than min pods running"        scale/scale.go:90  */
#define KSA_EV_UP_COOL 0x10      /* "Waiting for cool down, skipping scale up"   main.go:53 */
#define KSA_EV_DOWN_COOL 0x20    /* "Waiting for cool down, skipping scale down" main.go:67 */
/* Only the *_faults entry points produce these three (the plain ones replay a world without errors): */
#define KSA_EV_UP_FAILED 0x40    /* "Failed scaling up"    main.go:58 (tick ends, no stamp)           */
#define KSA_EV_DOWN_FAILED 0x80  /* "Failed scaling down"  main.go:72 (no stamp)                      */
#define KSA_EV_POLL_FAILED 0xC0  /* "Failed to get SQS messages" main.go:45: the event byte is exactly */
                                 /* this value (both failure bits: neither call could be attempted)   */

/* ---- failure mask (input of the *_faults entry points): what went wrong in a (tick, trace) ----- */
#define KSA_FAIL_POLL 0x01 /* sqs.NumMessages returned an error  main.go:43-47, sqs/sqs.go:51-54,59-61 */
#define KSA_FAIL_UP 0x02   /* p.ScaleUp returned an error         main.go:57-60, scale/scale.go:55-58,72-75 */
#define KSA_FAIL_DOWN 0x04 /* p.ScaleDown returned an error       main.go:71-74, scale/scale.go:82-85,100-103 */

/* ---- attribute axis: order of sqs.DefaultAttributeNames, sqs/sqs.go:28-32 ---------------- */
#define KSA_ATTR_MESSAGES 0x1u             /* ApproximateNumberOfMessages           */
#define KSA_ATTR_MESSAGES_DELAYED 0x2u     /* ApproximateNumberOfMessagesDelayed    */
#define KSA_ATTR_MESSAGES_NOT_VISIBLE 0x4u /* ApproximateNumberOfMessagesNotVisible */
#define KSA_MAX_ATTRS 3

/*
 * The policy record: the flag globals of main.go:16-25 (defaults main.go:83-91) plus the fields
 * of scale.PodAutoScaler (scale/scale.go:21-29). Pod counts are carried as Go `int` (int64) and
 * truncated with int32(...) inside the library exactly where scale.go does (:62,66-68,89,94-96).
 * Durations are Go time.Duration (int64 nanoseconds) and are converted to ticks inside the
 * library: cool_ticks = max(0, ceil(cool/poll)).
 */
typedef struct ksa_policy {
  int64_t scale_up_messages;       /* --scale-up-messages    main.go:86  (n >= this  => up)   */
  int64_t scale_down_messages;     /* --scale-down-messages  main.go:87  (n <= this  => down) */
  int64_t max_pods;                /* --max-pods             main.go:90                       */
  int64_t min_pods;                /* --min-pods             main.go:91                       */
  int64_t scale_up_pods;           /* --scale-up-pods        main.go:88                       */
  int64_t scale_down_pods;         /* --scale-down-pods      main.go:89                       */
  int64_t poll_period_ns;          /* --poll-period          main.go:83  must be > 0          */
  int64_t scale_up_cool_down_ns;   /* --scale-up-cool-down   main.go:85                       */
  int64_t scale_down_cool_down_ns; /* --scale-down-cool-down main.go:84                       */
  uint32_t attr_mask;              /* --attribute-names      main.go:93,103-110 (KSA_ATTR_*)  */
  uint32_t reserved;               /* must be 0                                               */
} ksa_policy;

/*
 * Per-trace carry state (structure of arrays, device memory, N elements each). It is what the
 * reference keeps between ticks: Spec.Replicas in the cluster (scale/scale.go:60,87) and the two
 * stamps lastScaleUpTime / lastScaleDownTime (main.go:37-38,62,76), the stamps expressed as the
 * absolute tick index at which they were taken (0 = process start). Replay calls read it,
 * advance it by T ticks and write it back, so a trace can be replayed in any number of chunks.
 */
typedef struct ksa_state {
  int32_t* replicas_dev;       /* [N] current replica count                          */
  int64_t* last_up_tick_dev;   /* [N] tick of the last lastScaleUpTime stamp, >= 0   */
  int64_t* last_down_tick_dev; /* [N] tick of the last lastScaleDownTime stamp, >= 0 */
} ksa_state;

/*
 * Per-trace summary (64 bytes). Counters accumulate across chunked calls when the caller does not
 * clear the buffer in between (ksa_summary_clear starts a fresh accumulation).
 *  hash: order-sensitive fingerprint of the (replicas, event) trajectory, two 32-bit lanes
 *        updated once per tick from h0 = 0 (hash = hi << 32 | lo):
 *          lo <- lo * 0x9E3779B1 + (uint32)replicas
 *          hi <- hi * 0x85EBCA6B + wide(event)
 *        wide(event) = the event byte with bit k (k = 0..5) moved to bit 5k and the two failure
 *        bits (6, 7) moved to bits 30, 31 (KSA_EV_WIDE below). Each lane is a polynomial in an
 *        odd multiplier over Z/2^32: any single-tick difference changes it. One integer
 *        multiply-add per lane and tick on the GPU.
 */
#define KSA_EV_WIDE(ev)                                                                        \
  (((((uint32_t)(ev)) & 7u) * 0x111u & 0x421u) | (((((uint32_t)(ev)) >> 3 & 7u) * 0x111u & 0x421u) << 15) | \
   ((((uint32_t)(ev)) >> 6) << 30))
typedef struct ksa_summary {
  int32_t final_replicas;
  int32_t min_replicas; /* over ticks replayed so far (INT32_MAX when none) */
  int32_t max_replicas; /* (INT32_MIN when none)                            */
  uint32_t up_applied;
  uint32_t down_applied;
  uint32_t up_noop;
  uint32_t down_noop;
  uint32_t up_cool;
  uint32_t down_cool;
  uint32_t poll_failed; /* the three failure counters stay 0 outside the *_faults entry points */
  uint32_t up_failed;
  uint32_t down_failed;
  int64_t sum_replicas; /* wrapping sum of the post-tick replica counts     */
  uint64_t hash;
} ksa_summary;

/* ---- synthetic trace generator (integer-only, counter-based; include/ksa_tracegen.h) ----- */
#define KSA_GEN_BANDS 0       /* i.i.d. three-band mixture around the thresholds          */
#define KSA_GEN_WALK 1        /* bounded integer random walk (sequential per trace)       */
#define KSA_GEN_ADVERSARIAL 2 /* threshold +-1, extremes, negatives, overflowing sums     */

typedef struct ksa_gen_params {
  uint64_t seed;
  int32_t family;        /* KSA_GEN_*                                                        */
  int32_t reserved;      /* must be 0                                                        */
  int64_t low_max;       /* BANDS: low band [0, low_max]            (40 % of entries)       */
  int64_t high_min;      /* BANDS: middle band (low_max, high_min)  (30 %)                  */
  int64_t high_max;      /* BANDS: high band [high_min, high_max]   (30 %); WALK: upper wall */
  int64_t walk_step;     /* WALK: step uniform in [-walk_step, +walk_step], reflect at walls */
  uint32_t overflow_log2; /* A==3 only: 1 entry in 2^overflow_log2 becomes a triple whose     */
                          /* int64 sum wraps; 0 = never                                       */
  uint32_t reserved2;
} ksa_gen_params;

/* ---- library / device ------------------------------------------------------------------- */

/* ABI version of the loaded library (== KSA_ABI_VERSION of the header it was built from). */
int ksa_abi_version(void);

/* Thread-local message describing the last non-zero return on this thread ("" if none). */
const char* ksa_last_error(void);

/* cudaSetDevice for callers that do not link the CUDA runtime themselves (Go, ctypes). */
int ksa_set_device(int device);

/* Number of visible CUDA devices, or a negative KSA_E* code. */
int ksa_device_count(void);

/* Convenience allocators for hosts without a CUDA binding. Not used by the replay calls. */
int ksa_dev_alloc(void** ptr_dev, size_t bytes);
int ksa_dev_free(void* ptr_dev);
int ksa_host_alloc(void** ptr_host, size_t bytes); /* pinned (page-locked) host memory */
int ksa_host_free(void* ptr_host);
/* Page-lock memory the caller already has (cudaHostRegister) so that copies to and from it run at
 * the link's rate and asynchronously. The case it is here for: the processes of one box (one per
 * GPU) map one shared segment, each registers its own slice and writes its shard's summaries
 * straight into it: the host-side gather of a sharded replay without a collective. ptr_host and
 * bytes should be multiples of the page size. Undo with ksa_host_unregister before unmapping. */
int ksa_host_register(void* ptr_host, size_t bytes);
int ksa_host_unregister(void* ptr_host);
int ksa_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream);
int ksa_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream);
int ksa_stream_sync(void* stream);

/* ---- policy helpers (host only) ---------------------------------------------------------- */

/* Fill *p with the reference's flag defaults, main.go:83-93 (poll 5 s, down-cool 30 s, up-cool
 * 10 s, up 100, down 10, steps 1/1, max 5, min 1, all three attributes). */
int ksa_policy_default(ksa_policy* p);

/* Cool-downs in ticks, as the kernels use them: max(0, ceil(cool/poll)). Replaces the wall-clock
 * test lastScaleXTime.Add(cool).After(time.Now()) of main.go:52,66. */
int ksa_policy_cool_ticks(const ksa_policy* p, int64_t* up_ticks, int64_t* down_ticks);

/* ---- the hot path ------------------------------------------------------------------------ */

/*
 * Replay T ticks of N independent traces, full trajectory out.
 * Replaces, per (trace, tick): NumMessages' attribute sum (sqs/sqs.go:56-64), Run's threshold and
 * cool-down tests and stamps (main.go:51-55,62,65-69,76), ScaleUp (scale/scale.go:60-69) and
 * ScaleDown (scale/scale.go:87-97).
 *   depth_dev    [T][A][ld] int64, A = number of attribute planes stored (1..3); the planes are
 *                the attributes selected by policy->attr_mask, lowest bit first. The kernel sums
 *                all A planes; popcount(attr_mask) must equal A.
 *   state        in/out carry state (see ksa_state)
 *   replicas_dev [T][ld] int32  replica count after each tick     (may be NULL: not written)
 *   events_dev   [T][ld] uint8  KSA_EV_* bits of each tick        (may be NULL: not written)
 * Limits: 1 <= T <= 2^30 - 64 per call (chunk longer traces: the carry state makes that exact),
 * 1 <= N <= 2^31-1, N <= ld.
 * Few traces, many ticks (N <= 49152, or 151552 < N <= 196608; T >= 4096, both outputs wanted): one thread per trace cannot
 * fill the GPU, so the library cuts the time axis into chunks that are replayed in parallel from
 * guessed states and then stitched exactly (csrc/ksa_timepar.cu; results are identical, the
 * trajectory and the carry state bit for bit). ksa_replay_summary and the two *_faults entry points
 * do the same below their own trace counts (32768, 16384 / as above): there the chunks' summary
 * records, including the fingerprint, are combined in tick order. Chunks whose guess never meets
 * the truth (a queue that stays saturated, idle or quiet) are worked out, shifted or replayed a second time,
 * and above 16384 traces, or with a failure mask, a 60 us probe decides first whether the depths suit the path at all; the
 * worst case found costs 1.06 x the plain kernels. That path keeps its per-chunk
 * records in fixed scratch arrays inside the library, so only one such call runs per GPU at a
 * time: a call that finds the scratch in use by a call on another stream, or a stream that is being
 * captured, takes the one-thread-per-trace kernels instead (same results, slower).
 */
int ksa_replay_full(const ksa_policy* policy, const int64_t* depth_dev, int64_t T, int32_t A,
                    int64_t N, int64_t ld, int64_t tick0, const ksa_state* state,
                    int32_t* replicas_dev, uint8_t* events_dev, void* stream);

/*
 * Same replay, no per-tick output: accumulates a ksa_summary per trace (read-modify-write of
 * summary_dev[N], so T-chunks accumulate). Same code path references as ksa_replay_full; the
 * counters are the reference's log lines counted (main.go:53,67; scale/scale.go:63,77,90,105).
 */
int ksa_replay_summary(const ksa_policy* policy, const int64_t* depth_dev, int64_t T, int32_t A,
                       int64_t N, int64_t ld, int64_t tick0, const ksa_state* state,
                       ksa_summary* summary_dev, void* stream);

/* Reset N summaries to the empty accumulation (counters 0, hash 0, min/max sentinels). */
int ksa_summary_clear(ksa_summary* summary_dev, int64_t N, void* stream);

/*
 * The same two replays in a world where calls fail: fail_dev [T][ld] uint8 carries KSA_FAIL_* bits
 * per (tick, trace) and reproduces the three error paths of Run, each of which `continue`s without
 * touching the stamps or the replicas:
 *   KSA_FAIL_POLL  NumMessages failed (main.go:43-47): the tick ends, event == KSA_EV_POLL_FAILED.
 *   KSA_FAIL_UP    ScaleUp failed (main.go:57-60): only matters if the tick reaches the call
 *                  (threshold met, cool-down over); no stamp, the tick ends (the down branch is
 *                  skipped), event bit KSA_EV_UP_FAILED.
 *   KSA_FAIL_DOWN  ScaleDown failed (main.go:71-74): likewise; an up-scale earlier in the same tick
 *                  stands; event bit KSA_EV_DOWN_FAILED.
 * KSA_FAIL_UP / KSA_FAIL_DOWN model a failure of the Get at the top of ScaleUp / ScaleDown
 * (scale/scale.go:55-58, 82-85): the call fails whether or not the replica count is at its limit.
 * A failing Update (scale/scale.go:72-75, 100-103) is the same thing away from the limit and cannot
 * happen at the limit (scale/scale.go:62-65, 89-92 return before it), so a recording that
 * distinguishes the two should not set the bit for an Update failure at the limit.
 * Algorithmic traffic: +1 byte per decision.
 */
int ksa_replay_full_faults(const ksa_policy* policy, const int64_t* depth_dev,
                           const uint8_t* fail_dev, int64_t T, int32_t A, int64_t N, int64_t ld,
                           int64_t tick0, const ksa_state* state, int32_t* replicas_dev,
                           uint8_t* events_dev, void* stream);
int ksa_replay_summary_faults(const ksa_policy* policy, const int64_t* depth_dev,
                              const uint8_t* fail_dev, int64_t T, int32_t A, int64_t N, int64_t ld,
                              int64_t tick0, const ksa_state* state, ksa_summary* summary_dev,
                              void* stream);

/*
 * Synthetic failure mask for ticks tick0+1..tick0+T of traces trace0..trace0+N-1: each KSA_FAIL_*
 * bit is set independently with probability 2^-poll_log2 / 2^-up_log2 / 2^-down_log2 (0 = never),
 * from the same counter-based hash as ksa_gen_traces (values depend only on seed, global trace id
 * and absolute tick).
 */
int ksa_gen_faults(uint64_t seed, uint32_t poll_log2, uint32_t up_log2, uint32_t down_log2,
                   int64_t tick0, int64_t T, int64_t trace0, int64_t N, int64_t ld,
                   uint8_t* fail_dev, void* stream);

/*
 * Policy sweep: replay the same N traces under P policies (the parameter grid of main.go:83-91),
 * summary per (policy, trace). Every policy starts every trace from replicas0_dev[n] with both
 * stamps at tick 0 (a fresh process, main.go:37-38). All policies must share A / attr_mask.
 *   policies      [P] host array
 *   policies_dev  [P] device scratch for the kernel-side policy table, sizeof(ksa_policy_k)=48 B
 *                 each (caller-allocated: P * 48 bytes)
 *   summary_dev   [P][N] ksa_summary, overwritten
 * The table is written by small kernels on `stream` that carry the converted records in their
 * parameters (no copy from a host buffer is left pending when the call returns). Policies are
 * replayed in register tiles of 4 consecutive policies; a tile whose policies share both message
 * thresholds (the usual layout of a parameter grid: thresholds vary slowest) runs fastest.
 */
int ksa_replay_sweep(const ksa_policy* policies, int32_t P, void* policies_dev,
                     const int64_t* depth_dev, int64_t T, int32_t A, int64_t N, int64_t ld,
                     const int32_t* replicas0_dev, ksa_summary* summary_dev, void* stream);

/*
 * The same sweep for callers that cannot allocate. (No counterpart in the reference: it has no
 * sweep, and no device memory to manage.)
 *
 * Who needs it: a host that embeds this library next to a framework which has already claimed the
 * GPU's memory for its own pool, so that a further cudaMalloc may simply fail; a Go host calling
 * through cgo with no CUDA binding of its own, for which "allocate 48*P bytes on the device and
 * keep them alive until the stream has drained" is one more lifetime to get wrong; and a caller
 * that re-runs sweeps from a latency-sensitive loop, where cudaMalloc and cudaFree are
 * unwelcome because they can stall the whole device while they run.
 *
 * What differs from ksa_replay_sweep: there is no policies_dev argument. The kernel-side policy
 * table lives in a fixed array inside the library's device image, KSA_SWEEP_STATIC_MAX_POLICIES
 * records of 48 bytes (192 KB), present on every GPU that loads the library until the process
 * exits. The call makes no cudaMalloc / cudaFree and no host heap allocation. Inputs, outputs and
 * results are exactly those of ksa_replay_sweep.
 *
 *  - P > KSA_SWEEP_STATIC_MAX_POLICIES returns KSA_ELIMIT; split the grid into several calls.
 *  - The first call on each device looks up the array's device address (once, with that device
 *    current). Make that first call before starting any stream capture on the device.
 *  - All calls on one device share the one array. It keeps the previous call's records; every call
 *    rewrites records [0, P) (kernels on `stream`, ahead of the sweep kernel) and nothing beyond
 *    P-1 is read, so stale contents are harmless. But two calls whose work overlaps on the same GPU (different streams,
 *    or different host threads) race on the table: run them on one stream, or wait for the first
 *    to finish. Calls on different GPUs do not interfere.
 */
#define KSA_SWEEP_STATIC_MAX_POLICIES 4096
int ksa_replay_sweep_static(const ksa_policy* policies, int32_t P, const int64_t* depth_dev,
                            int64_t T, int32_t A, int64_t N, int64_t ld,
                            const int32_t* replicas0_dev, ksa_summary* summary_dev, void* stream);

/*
 * Sweep reduction: score the P x N summaries of a sweep and pick policies. EXTENSION: the reference
 * has no cost model (it only has the tunables, main.go:83-91); the cost below is a plain linear
 * form over what the summaries count, in wrapping int64:
 *   cost(p, n) = w_pod_ticks * sum_replicas            (capacity paid for: pods x ticks)
 *              + w_saturated * up_noop                 (ticks over the up threshold while at max)
 *              + w_churn     * (up_applied + down_applied)
 *              + w_blocked   * (up_cool + down_cool)   (ticks a wanted scale was held by a cool-down)
 * Outputs (any may be NULL):
 *   policy_cost_dev [P] int64   sum over the N traces of cost(p, n); ACCUMULATED into, so that
 *                               trace shards (chunks, GPUs) add up: zero it before the first call
 *   best_policy_dev [N] int32   argmin over p of cost(p, n), lowest p on ties
 *   best_cost_dev   [N] int64   that minimum
 * One pass over the summaries (64 B per (policy, trace), HBM-bound). With traces sharded over GPUs
 * the per-policy totals are the one quantity that needs a cross-GPU step: an all-reduce (sum) of
 * the [P] vector, done by the host (kube_sqs_autoscaler_b200.shard.allreduce_policy_costs).
 */
typedef struct ksa_cost {
  int64_t w_pod_ticks;
  int64_t w_saturated;
  int64_t w_churn;
  int64_t w_blocked;
} ksa_cost;
int ksa_sweep_reduce(const ksa_summary* summary_dev, int32_t P, int64_t N, const ksa_cost* cost,
                     int64_t* policy_cost_dev, int32_t* best_policy_dev, int64_t* best_cost_dev,
                     void* stream);

/*
 * Fingerprint an existing full trajectory: hash_dev[n] = the ksa_summary.hash the summary replay
 * would produce for trace n over the same T ticks (h0 = hash_dev[n] on entry when chain != 0, else
 * 0). Used to compare 100+ GB trajectories without moving them to the host.
 */
int ksa_traj_hash(const int32_t* replicas_dev, const uint8_t* events_dev, int64_t T, int64_t N,
                  int64_t ld, int32_t chain, uint64_t* hash_dev, void* stream);

/* ---- inputs ------------------------------------------------------------------------------ */

/*
 * Write synthetic depths for ticks tick0+1..tick0+T of traces trace0..trace0+N-1 into
 * depth_dev [T][A][ld]. Values depend only on (seed, family, attribute, global trace id, absolute
 * tick), so any sharding / chunking produces the same bits. WALK needs walk_dev[N] (in/out, the
 * walk position after tick0; ksa_gen_walk_init fills the tick-0 position).
 */
int ksa_gen_traces(const ksa_gen_params* g, int64_t tick0, int64_t T, int32_t A, int64_t trace0,
                   int64_t N, int64_t ld, int64_t* depth_dev, int64_t* walk_dev, void* stream);
int ksa_gen_walk_init(const ksa_gen_params* g, int64_t trace0, int64_t N, int64_t* walk_dev,
                      void* stream);

/* Fresh carry state: replicas[n] = lo + hash(seed, trace0+n) mod (hi-lo+1), both stamps 0. */
int ksa_state_init(uint64_t seed, int64_t trace0, int64_t N, int32_t lo, int32_t hi,
                   const ksa_state* state, void* stream);

/* ---- ingest: recorded depths in the formats they arrive in ------------------------------- */

/*
 * Recorded queue-depth series usually arrive trace-major: one queue's history at a time
 * ([N][T], e.g. one metrics export per queue). The replay wants tick-major [T][A][ld]. This is a
 * tiled transpose through shared memory that also widens the element type.
 *   src_dev   [N][src_ld] elements of src_bytes (8 = int64, 4 = int32, sign-extended), src_ld >= T
 *   depth_dev [T][A][ld] int64; plane `a` (0 <= a < A) receives depth[(t*A + a)*ld + n] = src[n][t]
 * Replaces nothing in the reference (its only source is one live GetQueueAttributes per poll,
 * sqs/sqs.go:51); it is the step immediately before the path when replaying recorded data.
 * Traffic: src_bytes read + 8 written per element.
 */
int ksa_ingest_transpose(const void* src_dev, int32_t src_bytes, int64_t src_ld, int64_t T,
                         int64_t N, int32_t A, int32_t a, int64_t ld, int64_t* depth_dev,
                         void* stream);

/*
 * Decode SQS attribute values as the reference does: strconv.Atoi on a decimal string
 * (sqs/sqs.go:58). text_dev holds `count` fixed-width records of `width` bytes (1..32), each the
 * string followed by NUL padding (a record without NUL uses all `width` bytes).
 * Atoi semantics (Go 1.15 strconv): optional '+' or '-', then one or more ASCII digits, nothing
 * else; the value must fit a Go int (int64). Anything else is an error, for which the reference
 * returns from NumMessages with err != nil and Run skips the tick (sqs/sqs.go:59-61, main.go:44-47):
 *   values_dev[i] = the number, or 0 on error
 *   fail_dev[i]   = 0, or KSA_FAIL_POLL on error   (feeds the *_faults replays; may be NULL)
 * Traffic: width read + 9 written per record.
 */
int ksa_ingest_decimal(const char* text_dev, int32_t width, int64_t count, int64_t* values_dev,
                       uint8_t* fail_dev, void* stream);

/*
 * The failure bytes of ksa_ingest_decimal come out in the order of the records (trace-major when
 * the strings were); the *_faults replays want them tick-major:
 *   fail_dev[t*ld + n] = src_dev[n*src_ld + t]            (accumulate == 0)
 *   fail_dev[t*ld + n] |= src_dev[n*src_ld + t]           (accumulate != 0)
 * With several attributes, call it once per attribute with accumulate != 0 after the first: a
 * failed Atoi on any attribute fails the whole poll (sqs/sqs.go:57-61).
 */
int ksa_ingest_transpose_mask(const uint8_t* src_dev, int64_t src_ld, int64_t T, int64_t N, int64_t ld,
                              uint8_t* fail_dev, int32_t accumulate, void* stream);

/* ---- host-buffer pipeline (the call a Go / Python host makes with its own slices) --------- */

typedef struct ksa_pipeline ksa_pipeline; /* opaque; owns device staging + 3 streams */

/*
 * Create a streaming pipeline on `device` for up to N traces x A attributes, moving chunk_ticks
 * ticks per stage (double-buffered: 2 * chunk_ticks * N * (8A+6) bytes of device memory).
 */
int ksa_pipeline_create(ksa_pipeline** out, int device, int64_t N, int32_t A, int64_t chunk_ticks);
int ksa_pipeline_destroy(ksa_pipeline* p);

/*
 * Replay with HOST buffers: depth_host [T][A][N] int64 in, replicas_host [T][N] int32 and
 * events_host [T][N] uint8 out (either may be NULL), state_*_host [N] in/out (tight rows,
 * ld == N). H2D copy, kernel and D2H copy of consecutive chunks overlap on three streams.
 * Buffers should be pinned (ksa_host_alloc) for full PCIe rate; pageable memory works but is
 * slower. Synchronous: returns when the results are in the host buffers.
 * If summary_host != NULL it receives [N] ksa_summary for the T ticks (fresh accumulation).
 */
int ksa_pipeline_replay(ksa_pipeline* p, const ksa_policy* policy, const int64_t* depth_host,
                        int64_t T, int64_t N, int64_t tick0, int32_t* state_replicas_host,
                        int64_t* state_last_up_host, int64_t* state_last_down_host,
                        int32_t* replicas_host, uint8_t* events_host, ksa_summary* summary_host);

/*
 * The same with a failure mask: fail_host [T][N] uint8 of KSA_FAIL_* bits (tight rows), staged
 * with the depths (one more byte per decision host -> device) and replayed by
 * ksa_replay_full_faults / ksa_replay_summary_faults, i.e. Run's three error paths
 * (main.go:44-47, 57-60, 71-74). fail_host must not be NULL.
 */
int ksa_pipeline_replay_faults(ksa_pipeline* p, const ksa_policy* policy, const int64_t* depth_host,
                               const uint8_t* fail_host, int64_t T, int64_t N, int64_t tick0,
                               int32_t* state_replicas_host, int64_t* state_last_up_host,
                               int64_t* state_last_down_host, int32_t* replicas_host,
                               uint8_t* events_host, ksa_summary* summary_host);

/* ---- tuning / introspection -------------------------------------------------------------- */

/* Select a kernel variant for ksa_replay_full / ksa_replay_summary on the calling thread
 * (0 = library default, -2 = the scalar kernel, > 0 = an entry of the variant table). Used by the
 * tuning sweep in tools/; results never depend on the variant. Returns KSA_EINVAL if unknown. */
int ksa_set_variant(int32_t full_variant, int32_t summary_variant);

/* Name of a variant ("V4_D4_H1_B128x7": traces per thread, prefetch depth, load hint, block x
 * min-blocks-per-SM), or NULL if this build of the library has no such id. mode 0 = full, 1 =
 * summary, 2 / 3 = the same for the *_faults entry points (ids of KSA_FAULTS_VARIANT), 4 = launch
 * shapes of the sweep (ids of KSA_SWEEP_VARIANT); id 0 = the scalar kernel, id -1 = the library
 * default. The product build carries the kernels the library selects by itself; the tuning build
 * (-DKSA_TUNING_VARIANTS, `python __graft_entry__.py --tuning`) adds every point that was measured. */
const char* ksa_variant_name(int32_t mode, int32_t id);

/* Number of kernel launches issued by this library on the calling thread since the last call
 * (returns the count and resets it). bench.py reports it as gpu_launches. */
int64_t ksa_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* KSA_REPLAY_H */
