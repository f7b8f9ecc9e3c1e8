/*
 * ksa_oracle.h — CPU restatement of the kube-sqs-autoscaler scaling policy. TEST INFRASTRUCTURE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * call this; the product (libksareplay) never links or loads it.
 *
 * Parity status: the reference is Go and there is no Go toolchain here, so the reference itself
 * cannot be run. This restatement is PINNED by the 11 known-answer tests the reference
 * holds for this path (tests/golden/reference_kats.json: scale/scale_test.go:14-83,
 * sqs/sqs_test.go:19-25, main_test.go:19-239) and cross-checked against an independent Python
 * restatement (oracle/policy_py.py). Behaviour the reference's tests do not exercise (overlapping
 * thresholds, replicas outside [min,max], wrapping arithmetic, non-multiple cool-downs) is
 * "parity unpinned by the reference": it rests on the two restatements agreeing.
 */
#ifndef KSA_ORACLE_H
#define KSA_ORACLE_H

#include <stdint.h>
#include "../include/ksa_replay.h"

#ifdef __cplusplus
extern "C" {
#endif

/* sqs/sqs.go:56-64 — sum of the selected attributes, Go `int` (wrapping int64). */
int64_t ksa_oracle_num_messages(const int64_t* attrs, int32_t A);

/* scale/scale.go:54-79 — returns the KSA_EV_* bit of the log line it would print. */
uint8_t ksa_oracle_scale_up(const ksa_policy* p, int32_t* replicas);
/* scale/scale.go:81-107 */
uint8_t ksa_oracle_scale_down(const ksa_policy* p, int32_t* replicas);

/*
 * main.go:35-80 with a virtual wall clock, one trace: the literal restatement. Time is kept in
 * nanoseconds (128-bit so that tick*poll cannot overflow): tick k's checks happen at k*poll, the
 * stamps are the instant of the tick that took them, the cool-down test is the reference's
 * `last.Add(cool).After(now)`. depth is [T][A] for this one trace (tight).
 * state: replicas, last_up_ns / last_down_ns as multiples of poll given in TICKS.
 * fail: [T] KSA_FAIL_* bits for this trace (the three error paths main.go:44-47,57-60,71-74), or NULL.
 */
void ksa_oracle_run_wallclock(const ksa_policy* p, const int64_t* depth, const uint8_t* fail,
                              int64_t T, int32_t A, int64_t tick0, int32_t* replicas_state,
                              int64_t* last_up_tick, int64_t* last_down_tick, int32_t* replicas_out,
                              uint8_t* events_out);

/* max(0, ceil(cool/poll)) for both directions. Returns 0, or -1 if poll <= 0. */
int ksa_oracle_cool_ticks(const ksa_policy* p, int64_t* up_ticks, int64_t* down_ticks);

/*
 * Batch replay in the integer-tick model, same array layout as ksa_replay_full (ksa_replay.h):
 * depth [T][A][ld], fail [T][ld] (NULL: no failures), replicas/events [T][ld] (either may be NULL),
 * state arrays [N] in/out,
 * summary [N] accumulated when non-NULL. `threads` host threads split the trace axis
 * (<= 0: one per online CPU). Returns the number of threads used, or -1 on bad arguments.
 */
int ksa_oracle_replay(const ksa_policy* p, const int64_t* depth, const uint8_t* fail, int64_t T,
                      int32_t A, int64_t N, int64_t ld, int64_t tick0, int32_t* state_replicas, int64_t* state_last_up,
                      int64_t* state_last_down, int32_t* replicas_out, uint8_t* events_out,
                      ksa_summary* summary, int threads);

/* Empty accumulation, as ksa_summary_clear. */
void ksa_oracle_summary_clear(ksa_summary* s, int64_t N);

/* Trajectory fingerprint of existing outputs, as ksa_traj_hash (chain: continue from hash[n]). */
void ksa_oracle_traj_hash(const int32_t* replicas, const uint8_t* events, int64_t T, int64_t N,
                          int64_t ld, int32_t chain, uint64_t* hash);

/* Host side of the synthetic generator (include/ksa_tracegen.h), same layout and meaning as
 * ksa_gen_traces / ksa_gen_walk_init / ksa_state_init. `threads` as above. */
int ksa_oracle_gen_traces(const ksa_gen_params* g, int64_t tick0, int64_t T, int32_t A,
                          int64_t trace0, int64_t N, int64_t ld, int64_t* depth, int64_t* walk,
                          int threads);
void ksa_oracle_gen_walk_init(const ksa_gen_params* g, int64_t trace0, int64_t N, int64_t* walk);
void ksa_oracle_state_init(uint64_t seed, int64_t trace0, int64_t N, int32_t lo, int32_t hi,
                           int32_t* replicas, int64_t* last_up, int64_t* last_down);

/* Host side of ksa_gen_faults. */
void ksa_oracle_gen_faults(uint64_t seed, uint32_t poll_log2, uint32_t up_log2, uint32_t down_log2,
                           int64_t tick0, int64_t T, int64_t trace0, int64_t N, int64_t ld,
                           uint8_t* fail);

/* strconv.Atoi over fixed-width NUL-padded records, as ksa_ingest_decimal (sqs/sqs.go:58-61).
 * Returns the number of records that failed to parse. */
int64_t ksa_oracle_atoi(const char* text, int32_t width, int64_t count, int64_t* values, uint8_t* fail);

int ksa_oracle_online_cpus(void);

#ifdef __cplusplus
}
#endif
#endif
